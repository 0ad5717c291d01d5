#!/usr/bin/env python
"""Where the fixed cost of a long-context (B=1) decode step goes: graph-replayed step time over KV length x split count,
so that host time is out of the picture.  Run plain for step times, or under
`ncu --metrics gpu__time_duration.sum --clock-control none` for per-kernel durations."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def main():
    import mini_flash_attention as mfa
    quick = "--quick" in sys.argv
    dt, Hq, Hkv, D, page = torch.bfloat16, 32, 8, 128, 256
    for KV in ([16384] if quick else [16384, 65536, 131072]):
        nb = KV // page
        q = torch.randn(1, 1, Hq, D, device="cuda", dtype=dt)
        kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        bt = torch.randperm(nb, device="cuda").reshape(1, nb).int()
        sl = torch.full((1,), KV, dtype=torch.int32, device="cuda")
        ideal_us = 2.0 * KV * Hkv * D * 2 / 6.5e12 * 1e6
        for splits in ([0, 18] if quick else [0, 1, 4, 9, 18, 37, 74, 128]):
            fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits)  # noqa: E731
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            for _ in range(5):
                g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(50):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / 50 * 1e3
            print(f"KV={KV:7d} splits={splits:3d}  step {us:7.1f} us  (HBM ideal {ideal_us:5.1f} us, {2.0 * KV * Hkv * D * 2 / us / 1e3:7.1f} GB/s)", flush=True)


if __name__ == "__main__":
    main()
