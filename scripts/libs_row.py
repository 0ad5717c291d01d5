#!/usr/bin/env python
"""Library rows on the same box (comparison only, never on the product path): the official flash-attn 2.8.3 wheel
(what the reference's own benchmark compares against, /root/reference/benchmark/compare_prefill.py:42-55,
compare_decode.py:42-56) and, where it imports, flashinfer's prefill.  Same timing method as scripts/sweep.py
(5 warm-ups + 20 timed calls, CUDA events).  One JSON line per case, appended to gpurun_out/libs.jsonl.
    python scripts/libs_row.py [--quick]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402


def timed(fn, warm=5, n=20):
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


def emit(rec):
    line = json.dumps(rec)
    print(line, flush=True)
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "libs.jsonl"), "a") as f:
            f.write(line + "\n")


def main():
    quick = "--quick" in sys.argv
    import mini_flash_attention as mfa
    x = torch.randn(2, 512, 4, 128, device="cuda", dtype=torch.bfloat16)
    mfa.flash_attn_func(x, x, x)   # our library is up before the comparison libraries are imported
    torch.cuda.synchronize()
    try:
        import flash_attn
        from flash_attn import flash_attn_func as fa_func, flash_attn_with_kvcache as fa_kvcache
        fa_ver = flash_attn.__version__
    except Exception as e:  # noqa: BLE001
        emit({"lib": "flash_attn", "error": f"{type(e).__name__}: {e}"[:200]})
        fa_func = fa_kvcache = None
        fa_ver = None
    fi = None
    try:
        import flashinfer
        fi = flashinfer
    except Exception as e:  # noqa: BLE001
        emit({"lib": "flashinfer", "error": f"{type(e).__name__}: {e}"[:200]})

    g = torch.Generator(device="cuda").manual_seed(0)
    B, H, D = 48, 24, 128
    dt = torch.bfloat16
    for S in ((4096,) if quick else (256, 512, 1024, 2048, 4096)):
        q, k, v = (torch.randn(B, S, H, D, device="cuda", dtype=dt, generator=g) for _ in range(3))
        for causal in (False, True):
            fl = 4.0 * B * H * D * (S * (S + 1) / 2 if causal else S * S)
            ms = timed(lambda: mfa.flash_attn_func(q, k, v, causal=causal))
            rec = {"case": "prefill", "dtype": "bfloat16", "S": S, "causal": causal, "ours_ms": round(ms, 4),
                   "ours_tflops": round(fl / ms / 1e9, 1)}
            if fa_func is not None:
                try:
                    o_fa = fa_func(q, k, v, causal=causal)
                    o_us = mfa.flash_attn_func(q, k, v, causal=causal)
                    rec["maxdiff_vs_flash_attn"] = float((o_fa[:2].float() - o_us[:2].float()).abs().max())
                    fms = timed(lambda: fa_func(q, k, v, causal=causal), 3, 10)
                    rec.update(flash_attn=fa_ver, flash_attn_ms=round(fms, 4), flash_attn_tflops=round(fl / fms / 1e9, 1))
                except Exception as e:  # noqa: BLE001
                    rec["flash_attn_error"] = f"{type(e).__name__}: {e}"[:200]
            if fi is not None and S == 4096:
                try:
                    # flashinfer ragged prefill over the batch (its Blackwell FMHA when the backend is available)
                    qo = torch.arange(0, (B + 1) * S, S, dtype=torch.int32, device="cuda")
                    ws = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
                    for backend in ("cutlass", "auto"):
                        try:
                            w = fi.BatchPrefillWithRaggedKVCacheWrapper(ws, "NHD", backend=backend)
                            w.plan(qo, qo, H, H, D, causal=causal, q_data_type=dt)
                            q2, k2, v2 = (t.view(B * S, H, D) for t in (q, k, v))
                            o_fi = w.run(q2, k2, v2)
                            fims = timed(lambda: w.run(q2, k2, v2), 3, 10)
                            rec[f"flashinfer_{backend}_ms"] = round(fims, 4)
                            rec[f"flashinfer_{backend}_tflops"] = round(fl / fims / 1e9, 1)
                            rec[f"maxdiff_vs_flashinfer_{backend}"] = float(
                                (o_fi.view(B, S, H, D)[:2].float() - mfa.flash_attn_func(q, k, v, causal=causal)[:2].float()).abs().max())
                        except Exception as e:  # noqa: BLE001
                            rec[f"flashinfer_{backend}_error"] = f"{type(e).__name__}: {e}"[:160]
                    del ws
                except Exception as e:  # noqa: BLE001
                    rec["flashinfer_error"] = f"{type(e).__name__}: {e}"[:200]
            emit(rec)
        del q, k, v
    # ---- decode, paged(256)
    page, dB, dH = 256, 24, 24
    for Hq, Hkv in ((24, 24), (32, 8)):
        for KV in ((8192,) if quick else (512, 1024, 2048, 4096, 8192)):
            per = KV // page
            nb = dB * per
            dq = torch.randn(dB, 1, Hq, D, device="cuda", dtype=dt, generator=g)
            kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt, generator=g)
            vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt, generator=g)
            bt = torch.randperm(nb, device="cuda", generator=g).reshape(dB, per).int()
            sl = torch.full((dB,), KV, dtype=torch.int32, device="cuda")
            nbytes = 2.0 * dB * Hkv * KV * D * 2 + 2.0 * dq.numel() * 2
            ms = timed(lambda: mfa.flash_attn_with_kvcache(dq, kc, vc, sl, bt, 0), 10, 50)
            rec = {"case": "decode", "Hq": Hq, "Hkv": Hkv, "KV": KV, "ours_us": round(ms * 1e3, 2), "ours_GBps": round(nbytes / ms / 1e6, 1)}
            if fa_kvcache is not None:
                try:
                    o_fa = fa_kvcache(dq, kc, vc, cache_seqlens=sl, block_table=bt, causal=True)
                    o_us = mfa.flash_attn_with_kvcache(dq, kc, vc, sl, bt, 0)
                    rec["maxdiff_vs_flash_attn"] = float((o_fa.float() - o_us.float()).abs().max())
                    fms = timed(lambda: fa_kvcache(dq, kc, vc, cache_seqlens=sl, block_table=bt, causal=True), 10, 50)
                    rec.update(flash_attn_us=round(fms * 1e3, 2), flash_attn_GBps=round(nbytes / fms / 1e6, 1))
                except Exception as e:  # noqa: BLE001
                    rec["flash_attn_error"] = f"{type(e).__name__}: {e}"[:200]
            emit(rec)


if __name__ == "__main__":
    main()
