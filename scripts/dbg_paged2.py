import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import mini_flash_attention as mfa
torch.manual_seed(0)
D, page, Sk, Sq = 128, int(sys.argv[1]) if len(sys.argv) > 1 else 8, 256, 50
nb = Sk // page
q = torch.randn(Sq, 1, D, device="cuda", dtype=torch.float16)
kc = torch.randn(nb, page, 1, D, device="cuda", dtype=torch.float16)
vc = torch.randn(nb, page, 1, D, device="cuda", dtype=torch.float16)
perm = torch.arange(nb, device="cuda")  # identity table first
for name, table in (("identity", perm), ("reversed", perm.flip(0))):
    bt = table.reshape(1, nb).int().contiguous()
    cu_q = torch.tensor([0, Sq], dtype=torch.int32, device="cuda"); cu_k = torch.tensor([0, Sk], dtype=torch.int32, device="cuda")
    out = mfa.flash_attn_varlen_func(q, kc, vc, cu_q, cu_k, Sq, Sk, causal=False, block_table=bt).float()[:, 0]
    K = kc[table.long()].reshape(Sk, D).float(); V = vc[table.long()].reshape(Sk, D).float()
    s = (q[:, 0].float() @ K.t()) / D ** 0.5
    def att(mask):
        ss = s.clone(); ss[:, ~mask] = -1e30
        return torch.softmax(ss, -1) @ V
    full = torch.ones(Sk, dtype=torch.bool, device="cuda")
    print(name, "err full", (out - att(full)).abs().max().item())
    for lo in range(0, Sk, 64):
        m = full.clone(); m[lo:lo + 64] = False
        print("   without keys", lo, lo + 64, (out - att(m)).abs().max().item())
    # V-only check: uniform attention (q = 0) isolates V loading; 
q0 = torch.zeros_like(q)
bt = perm.reshape(1, nb).int().contiguous()
out = mfa.flash_attn_varlen_func(q0, kc, vc, cu_q, cu_k, Sq, Sk, causal=False, block_table=bt).float()[:, 0]
V = vc.reshape(Sk, D).float()
print("q=0 (mean of V) err", (out - V.mean(0, keepdim=True)).abs().max().item(), "err per 16-col group", [(out - V.mean(0, keepdim=True)).abs()[:, c:c + 16].max().item() for c in range(0, D, 16)])
for lo in range(0, Sk, 64):
    Vm = torch.cat([V[:lo], V[lo + 64:]]).sum(0, keepdim=True) / Sk
    print("   q=0 without V rows", lo, (out - Vm).abs().max().item())
