#!/usr/bin/env python
"""head_dim 160 / 192 prefill: the DPAD = 192 kernel against zero-padding to 256 (MFA_PREFILL_NO_D192=1 in the environment)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
print("mode:", "padded to 256" if os.environ.get("MFA_PREFILL_NO_D192") == "1" else "DPAD = 192")
for D in (160, 192):
    for causal in (False, True):
        B, S, H = 16, 4096, 16
        q, k, v = (torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16) for _ in range(3))
        for _ in range(3): mfa.flash_attn_func(q, k, v, causal=causal)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): mfa.flash_attn_func(q, k, v, causal=causal)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        fl = 4 * B * H * S * S * D * (0.5 if causal else 1.0)
        print(f"D={D} causal={causal} B{B} H{H} S{S}: {ms:.3f} ms = {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
