#!/usr/bin/env python
"""Kernel-tuning aid (GPU box): time the headline prefill with each alternative build of the library.
    python scripts/tune_prefill.py libA.so libB.so ...      (paths; 'default' = the in-tree lib)
Each lib runs in its own process (MFA_B200_LIB).  Prints ms, TFLOP/s, max-abs error vs the fp32 oracle on a
small case, and SM clock during the run."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

def child():
    import torch
    sys.path.insert(0, ROOT)
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    causal = os.environ.get("TUNE_CAUSAL") == "1"
    B, H, S, D = 48, 24, int(os.environ.get("TUNE_S", "4096")), 128
    dt = torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn(B, S, H, D, device="cuda", dtype=dt, generator=g) for _ in range(3))
    qs, ks, vs = q[:1, :1024, :2].contiguous(), k[:1, :1024, :2].contiguous(), v[:1, :1024, :2].contiguous()
    err = (mfa.flash_attn_func(qs, ks, vs, causal=causal).float() - R.attention_ref(qs, ks, vs, causal)).abs().max().item()
    for _ in range(5):
        mfa.flash_attn_func(q, k, v, causal=causal)
    torch.cuda.synchronize()
    best, tot = 1e9, 0.0
    n = 15
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); mfa.flash_attn_func(q, k, v, causal=causal); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1); best = min(best, ms); tot += ms
    fl = 4.0 * B * H * S * S * D * (0.5 if causal else 1.0)
    # the bench's method: launches back to back, one pair of events around all of them (two rounds, both reported)
    b2b = []
    for _ in range(2):
        nb = int(os.environ.get("TUNE_B2B", "20"))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(nb):
            mfa.flash_attn_func(q, k, v, causal=causal)
        e1.record(); torch.cuda.synchronize()
        b2b.append(round(fl / (e0.elapsed_time(e1) / nb) / 1e9, 1))
    clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader,nounits"], capture_output=True, text=True).stdout.strip()
    print("TUNE " + json.dumps({"lib": os.path.basename(os.environ.get("MFA_B200_LIB", "default")), "ms_avg": round(tot / n, 3), "ms_best": round(best, 3),
                                "tflops_avg": round(fl / (tot / n) / 1e9, 1), "tflops_best": round(fl / best / 1e9, 1), "tflops_b2b": b2b, "err": err, "clk_after": clk}))

if __name__ == "__main__":
    if os.environ.get("TUNE_CHILD"):
        child(); sys.exit(0)
    for lib in sys.argv[1:]:
        env = dict(os.environ, TUNE_CHILD="1")
        if lib != "default":
            env["MFA_B200_LIB"] = os.path.abspath(lib)
        r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=300)
        out = [l for l in r.stdout.splitlines() if l.startswith("TUNE ")]
        print(out[-1] if out else ("FAIL " + lib + " " + r.stderr[-400:]), flush=True)
