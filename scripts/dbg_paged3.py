import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
torch.manual_seed(0)
D, page, Sk, Sq = 128, int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 256, 50
nb = Sk // page
q = torch.randn(Sq, 1, D, device="cuda", dtype=torch.float16)
kc = torch.randn(nb, page, 1, D, device="cuda", dtype=torch.float16)
bt = torch.arange(nb, device="cuda").reshape(1, nb).int().contiguous()
cu_q = torch.tensor([0, Sq], dtype=torch.int32, device="cuda"); cu_k = torch.tensor([0, Sk], dtype=torch.int32, device="cuda")
K = kc.reshape(Sk, D).float()
Pref = torch.softmax((q[:, 0].float() @ K.t()) / D ** 0.5, -1)
for t in range(Sk // 128):
    V = torch.zeros(Sk, D, device="cuda", dtype=torch.float16)
    V[t * 128:(t + 1) * 128] = torch.eye(128, device="cuda", dtype=torch.float16)
    vc = V.reshape(nb, page, 1, D).contiguous()
    out = mfa.flash_attn_varlen_func(q, kc, vc, cu_q, cu_k, Sq, Sk, causal=False, block_table=bt).float()[:, 0]
    err = (out - Pref[:, t * 128:(t + 1) * 128]).abs()
    print("tile", t, "P err max", err.max().item(), "per 8-col group:", [round(err[:, c:c + 8].max().item(), 3) for c in range(0, 128, 8)])
