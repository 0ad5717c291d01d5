#!/usr/bin/env python
"""Sweep of BASELINE.json's configs on one GPU (run on the B200 box), ours next to the reference's own CUDA kernels
(oracle/_ref) on the same device:
  * prefill flash_attn_func d=128 B=48 H=24, S = 256 ... 4096, causal and non-causal, bf16 and fp16 (configs[1])
  * varlen continuous batching [128 ... 4096], Hq=32, Hkv=8, causal, packed and paged(256)             (configs[2])
  * decode q_len=1 d=128 B=24 H=24, KV = 512 ... 8192, paged(256), num_splits auto                      (configs[3])
Prints one JSON line per case (and appends to gpurun_out/sweep.jsonl).  CUDA events, 5 warm-ups + 20 timed calls."""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mini_flash_attention as mfa  # noqa: E402

PEAK_TF, PEAK_GB = 1628.4, 6547.8
p = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(p):
    d = json.load(open(p))
    PEAK_TF, PEAK_GB = d["bf16_tflops"], d["hbm_gbs"]


def load_ref():
    so = os.path.join(ROOT, "oracle", "_ref", "mini_flash_attention_ref_C.so")
    if not os.path.exists(so):
        return None
    try:
        spec = importlib.util.spec_from_file_location("mini_flash_attention_ref_C", so)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception:  # noqa: BLE001
        return None


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
        with open(os.path.join(d, "sweep.jsonl"), "a") as f:
            f.write(line + "\n")


def main():
    ref = load_ref()
    g = torch.Generator(device="cuda").manual_seed(0)
    B, H, D = 48, 24, 128
    for dt in (torch.bfloat16, torch.float16):
        for S in (256, 512, 1024, 2048, 4096):
            q, k, v = (torch.randn(B, S, H, D, device="cuda", dtype=dt, generator=g) for _ in range(3))
            for causal in (False, True):
                if dt == torch.float16 and S not in (1024, 4096):
                    continue
                fl = 4.0 * B * H * D * (S * (S + 1) / 2 if causal else S * S)
                ms = timed(lambda: mfa.flash_attn_func(q, k, v, causal=causal))
                rec = {"case": "prefill", "dtype": str(dt).split(".")[-1], "S": S, "causal": causal, "ms": round(ms, 4),
                       "tflops": round(fl / ms / 1e9, 1), "frac_of_measured_peak": round(fl / ms / 1e9 / PEAK_TF, 3),
                       "hbm_GBps": round(4.0 * B * S * H * D * 2 / ms / 1e6, 1)}
                if ref is not None:
                    rms = timed(lambda: ref.mini_flash_attention_forward(q, k, v, None, causal, -1, -1), 2, 5)
                    rec.update(ref_ms=round(rms, 4), ref_tflops=round(fl / rms / 1e9, 1), speedup=round(rms / ms, 2))
                emit(rec)
            del q, k, v
    # ---- varlen continuous batching, GQA, causal; packed and paged(256)
    seqlens = [128, 256, 512, 1024, 2048, 4096]
    Hq, Hkv = 32, 8
    tot = sum(seqlens)
    cu = torch.tensor([0] + list(torch.tensor(seqlens).cumsum(0)), dtype=torch.int32, device="cuda")
    q = torch.randn(tot, Hq, D, device="cuda", dtype=torch.bfloat16, generator=g)
    k = torch.randn(tot, Hkv, D, device="cuda", dtype=torch.bfloat16, generator=g)
    v = torch.randn(tot, Hkv, D, device="cuda", dtype=torch.bfloat16, generator=g)
    fl = sum(4.0 * Hq * D * s * (s + 1) / 2 for s in seqlens)
    ms = timed(lambda: mfa.flash_attn_varlen_func(q, k, v, cu, cu, max(seqlens), max(seqlens), causal=True))
    rec = {"case": "varlen packed", "seqlens": seqlens, "Hq": Hq, "Hkv": Hkv, "causal": True, "ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1)}
    if ref is not None:
        rms = timed(lambda: ref.mini_flash_attention_varlen_forward(q, k, v, cu, cu, max(seqlens), max(seqlens), True, -1, -1, None), 2, 5)
        rec.update(ref_ms=round(rms, 4), speedup=round(rms / ms, 2))
    emit(rec)
    page = 256
    nblk = sum(-(-s // page) for s in seqlens)
    perm = torch.randperm(nblk, device="cuda", generator=g)
    maxb = max(-(-s // page) for s in seqlens)
    bt = torch.zeros(len(seqlens), maxb, dtype=torch.int32, device="cuda")
    kc = torch.zeros(nblk, page, Hkv, D, device="cuda", dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    used = 0
    for i, s in enumerate(seqlens):
        nb = -(-s // page)
        ids = perm[used:used + nb]
        used += nb
        bt[i, :nb] = ids.int()
        ks = k[cu[i]:cu[i + 1]]
        vs = v[cu[i]:cu[i + 1]]
        pad = nb * page - s
        kc[ids] = torch.cat([ks, ks.new_zeros(pad, Hkv, D)]).view(nb, page, Hkv, D)
        vc[ids] = torch.cat([vs, vs.new_zeros(pad, Hkv, D)]).view(nb, page, Hkv, D)
    out_packed = mfa.flash_attn_varlen_func(q, k, v, cu, cu, max(seqlens), max(seqlens), causal=True)
    out_paged = mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, max(seqlens), max(seqlens), causal=True, block_table=bt)
    ms = timed(lambda: mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, max(seqlens), max(seqlens), causal=True, block_table=bt))
    rec = {"case": "varlen paged(256)", "seqlens": seqlens, "Hq": Hq, "Hkv": Hkv, "causal": True, "ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1),
           "paged_equals_packed_maxdiff": float((out_packed.float() - out_paged.float()).abs().max())}
    if ref is not None:
        rms = timed(lambda: ref.mini_flash_attention_varlen_forward(q, kc, vc, cu, cu, max(seqlens), max(seqlens), True, -1, -1, bt), 2, 5)
        rec.update(ref_ms=round(rms, 4), speedup=round(rms / ms, 2))
    emit(rec)
    del q, k, v, kc, vc
    # ---- decode
    dB, dH = 24, 24
    for Hkv_d in (24, 8):
        for KV in (512, 1024, 2048, 4096, 8192):
            per = KV // page
            nb = dB * per
            dq = torch.randn(dB, 1, dH if Hkv_d == 24 else 32, D, device="cuda", dtype=torch.bfloat16, generator=g)
            kc = torch.randn(nb, page, Hkv_d, D, device="cuda", dtype=torch.bfloat16, generator=g)
            vc = torch.randn(nb, page, Hkv_d, D, device="cuda", dtype=torch.bfloat16, generator=g)
            bt = torch.randperm(nb, device="cuda", generator=g).reshape(dB, per).int()
            sl = torch.full((dB,), KV, dtype=torch.int32, device="cuda")
            nbytes = 2.0 * dB * Hkv_d * KV * D * 2 + 2.0 * dq.numel() * 2
            ms = timed(lambda: mfa.flash_attn_with_kvcache(dq, kc, vc, sl, bt, 0), 10, 50)
            rec = {"case": "decode", "Hq": dq.shape[2], "Hkv": Hkv_d, "KV": KV, "us": round(ms * 1e3, 2), "GBps": round(nbytes / ms / 1e6, 1),
                   "frac_of_measured_hbm": round(nbytes / ms / 1e6 / PEAK_GB, 3)}
            if ref is not None:
                rms = timed(lambda: ref.mini_flash_attention_with_kvcache(dq, kc, vc, sl, bt, False, 0), 5, 20)
                rec.update(ref_us=round(rms * 1e3, 2), ref_GBps=round(Hkv_d and nbytes / rms / 1e6, 1), speedup=round(rms / ms, 2))
            emit(rec)


if __name__ == "__main__":
    main()
