#!/usr/bin/env python
"""Tuning aid (GPU box): decode B=24 D=128 paged(256), MHA 24/24 and GQA 32/8, KV 512 ... 8192 (BASELINE configs[3]),
bf16 and e4m3 caches, with the library's environment switches, one child process per setting:
    python scripts/decode_ab.py [ENV=VAL ...]      e.g.  python scripts/decode_ab.py MFA_DECODE_NONPERSISTENT=1
Eager calls (50 back to back, CUDA events) and CUDA-graph replay of one call; GB/s of algorithmic bytes."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    import torch
    sys.path.insert(0, ROOT)
    import mini_flash_attention as mfa
    tag = os.environ.get("AB_TAG", "default")
    g = torch.Generator(device="cuda").manual_seed(0)
    B, D, page = 24, 128, 256

    def timed(fn, warm=10, n=50):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    for fp8 in (False, True):
        for Hq, Hkv in ((24, 24), (32, 8)):
            for KV in (512, 1024, 2048, 4096, 8192):
                nb = B * KV // page
                q = torch.randn(B, 1, Hq, D, device="cuda", dtype=torch.bfloat16, generator=g)
                kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=torch.bfloat16, generator=g)
                vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=torch.bfloat16, generator=g)
                kw = {}
                if fp8:
                    kc, vc = (kc * 0.25).to(torch.float8_e4m3fn), vc.to(torch.float8_e4m3fn)
                    kw = dict(k_scale=4.0, v_scale=1.0)
                bt = torch.randperm(nb, device="cuda", generator=g).reshape(B, KV // page).int()
                sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
                fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, **kw)  # noqa: E731
                ms = timed(fn)
                gr = torch.cuda.CUDAGraph()
                fn(); torch.cuda.synchronize()
                with torch.cuda.graph(gr):
                    fn()
                gms = timed(gr.replay)
                nbytes = 2.0 * B * Hkv * KV * D * kc.element_size() + 2.0 * q.numel() * 2
                print("AB " + json.dumps({"tag": tag, "cache": "e4m3" if fp8 else "bf16", "Hq": Hq, "Hkv": Hkv, "KV": KV, "us": round(ms * 1e3, 2),
                                          "GBps": round(nbytes / ms / 1e6, 1), "graph_us": round(gms * 1e3, 2),
                                          "graph_GBps": round(nbytes / gms / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    if os.environ.get("AB_CHILD"):
        child(); sys.exit(0)
    settings = [a for a in sys.argv[1:] if "=" in a]
    for env_extra, tag in ([({}, "default")] + [({s.split("=")[0]: s.split("=")[1]}, s) for s in settings]):
        env = dict(os.environ, AB_CHILD="1", AB_TAG=tag, **env_extra)
        r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=600)
        out = [l for l in r.stdout.splitlines() if l.startswith("AB ")]
        print("\n".join(out) if out else ("FAIL " + tag + " " + r.stderr[-600:]), flush=True)
