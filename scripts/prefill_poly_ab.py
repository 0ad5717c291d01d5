import sys, os, torch
sys.path.insert(0, "/root/repo")
import mini_flash_attention as mfa
S = int(sys.argv[1]); causal = sys.argv[2] == "1"
B,H,D=48,24,128
q,k,v=(torch.randn(B,S,H,D,device="cuda",dtype=torch.bfloat16) for _ in range(3))
for _ in range(10): mfa.flash_attn_func(q,k,v,causal=causal)
torch.cuda.synchronize()
n = max(20, int(200e-3 / (4.0*B*H*S*S*D/1.0e15)))
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n): mfa.flash_attn_func(q,k,v,causal=causal)
e1.record(); torch.cuda.synchronize()
ms=e0.elapsed_time(e1)/n
fl=4.0*B*H*S*S*D*(0.5 if causal else 1)
print(f"S={S} causal={causal} POLY={os.environ.get('MFA_PREFILL_POLY')} n={n} {ms:.3f} ms {fl/ms/1e9:.1f} TFLOP/s", flush=True)
