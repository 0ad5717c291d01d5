#!/usr/bin/env python
"""Bring-up diagnostics on a B200 box: runs each check in its own subprocess (a device trap poisons the CUDA
context) with a timeout, prints one line per check and writes gpurun_out/diag.json.

    python scripts/gpu_diag.py [name-substring ...]
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHECKS = {}


def check(fn):
    CHECKS[fn.__name__] = fn
    return fn


def _imports():
    import torch
    from mini_flash_attention_b200 import _lib
    return torch, _lib


@check
def selftest_qk_k64():
    return _selftest(0, 64, 128)


@check
def selftest_qk_k128():
    return _selftest(0, 128, 128)


@check
def selftest_pv_n64():
    return _selftest(1, 128, 64)


@check
def selftest_pv_n128():
    return _selftest(1, 128, 128)


@check
def selftest_pv_n256():
    return _selftest(1, 128, 256)


@check
def selftest_qk_bf16():
    return _selftest(0, 128, 128, bf16=True)


def _selftest(mode, kdim, ncols, bf16=False, lbo=0, sbo=0):
    torch, _lib = _imports()
    lib = _lib.load()
    torch.manual_seed(0)
    dt = torch.bfloat16 if bf16 else torch.float16
    if mode == 0:
        a = torch.randn(128, kdim, device="cuda", dtype=dt)
        b = torch.randn(128, kdim, device="cuda", dtype=dt)
        ref = a.float() @ b.float().t()
    else:
        a = torch.rand(128, 128, device="cuda", dtype=dt)
        b = torch.randn(128, ncols, device="cuda", dtype=dt)
        ref = a.float() @ b.float()
    d = torch.zeros(128, ncols if mode == 1 else 128, device="cuda", dtype=torch.float32)
    rc = lib.mfa_selftest_umma(mode, a.data_ptr(), b.data_ptr(), d.data_ptr(), kdim, ncols, int(bf16), lbo, sbo, None)
    torch.cuda.synchronize()
    err = (d - ref).abs().max().item()
    return {"rc": rc, "max_err": err, "ref_absmax": ref.abs().max().item(), "ok": rc == 0 and err < 0.05}


def _prefill_case(B, Sq, Sk, H, Hkv, D, causal, dtype="fp16"):
    torch, _ = _imports()
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    torch.manual_seed(0)
    dt = torch.float16 if dtype == "fp16" else torch.bfloat16
    q = torch.randn(B, Sq, H, D, device="cuda", dtype=dt)
    k = torch.randn(B, Sk, Hkv, D, device="cuda", dtype=dt)
    v = torch.randn(B, Sk, Hkv, D, device="cuda", dtype=dt)
    out = mfa.flash_attn_func(q, k, v, causal=causal)
    torch.cuda.synchronize()
    ref = R.attention_ref(q, k, v, causal)
    err = (out.float() - ref).abs().max().item()
    tol = 2e-3 if dtype == "fp16" else 1e-2
    return {"max_err": err, "finite": bool(torch.isfinite(out).all()), "ok": err <= tol}


@check
def prefill_128_d128():
    return _prefill_case(1, 128, 128, 1, 1, 128, False)


@check
def prefill_256_d128():
    return _prefill_case(1, 256, 256, 2, 2, 128, False)


@check
def prefill_512_d64():
    return _prefill_case(2, 512, 512, 4, 4, 64, False)


@check
def prefill_1024_causal():
    return _prefill_case(2, 1024, 1024, 8, 8, 128, True)


@check
def prefill_odd_gqa():
    return _prefill_case(2, 333, 333, 8, 2, 128, True)


@check
def prefill_bf16():
    return _prefill_case(1, 512, 512, 4, 4, 128, False, "bf16")


@check
def prefill_d256():
    return _prefill_case(1, 300, 300, 2, 2, 256, True)


@check
def prefill_d96_d32():
    a = _prefill_case(1, 256, 256, 4, 4, 96, False)
    b = _prefill_case(1, 256, 256, 4, 4, 32, True)
    return {"d96": a, "d32": b, "ok": a["ok"] and b["ok"]}


@check
def prefill_4096():
    return _prefill_case(1, 4096, 4096, 2, 2, 128, False)


def _decode_case(B, H, Hkv, D, KV, page=None, splits=0, dtype="fp16", ragged=False):
    torch, _ = _imports()
    import mini_flash_attention as mfa
    from oracle import ref_attention as R
    torch.manual_seed(0)
    dt = torch.float16 if dtype == "fp16" else torch.bfloat16
    q = torch.randn(B, 1, H, D, device="cuda", dtype=dt)
    if ragged:
        lens = torch.randint(1, KV + 1, (B,), device="cuda", dtype=torch.int32)
        lens[0] = KV
    else:
        lens = torch.full((B,), KV, device="cuda", dtype=torch.int32)
    if page:
        per = (KV + page - 1) // page
        nb = B * per + 3
        kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        bt = torch.randperm(nb, device="cuda")[: B * per].reshape(B, per).int()
    else:
        kc = torch.randn(B, KV, Hkv, D, device="cuda", dtype=dt)
        vc = torch.randn(B, KV, Hkv, D, device="cuda", dtype=dt)
        bt = None
    out = mfa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, num_splits=splits)
    torch.cuda.synchronize()
    ref = R.kvcache_ref(q, kc, vc, lens, bt)
    err = (out.float() - ref).abs().max().item()
    tol = 2e-3 if dtype == "fp16" else 1e-2
    return {"max_err": err, "finite": bool(torch.isfinite(out).all()), "ok": err <= tol}


@check
def decode_contig_512():
    return _decode_case(2, 8, 8, 128, 512)


@check
def decode_paged256_gqa():
    return _decode_case(4, 8, 2, 128, 1000, page=256, ragged=True)


@check
def decode_paged16_split():
    return _decode_case(3, 8, 8, 64, 777, page=16, splits=3, ragged=True)


@check
def decode_d256_bf16():
    return _decode_case(2, 4, 4, 256, 300, page=64, dtype="bf16")


@check
def decode_auto_split_long():
    return _decode_case(1, 32, 8, 128, 16384, page=256)


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--run":
        name = sys.argv[2]
        try:
            res = CHECKS[name]()
        except Exception as e:  # noqa: BLE001
            res = {"ok": False, "exception": f"{type(e).__name__}: {e}"[:600]}
        print("DIAG_RESULT " + json.dumps(res))
        return
    filt = sys.argv[1:]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    results = {}
    for name in CHECKS:
        if filt and not any(f in name for f in filt):
            continue
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--run", name], capture_output=True,
                               text=True, timeout=180, cwd=ROOT)
            line = [l for l in r.stdout.splitlines() if l.startswith("DIAG_RESULT ")]
            res = json.loads(line[-1][len("DIAG_RESULT "):]) if line else {
                "ok": False, "rc": r.returncode, "stdout": r.stdout[-800:], "stderr": r.stderr[-1500:]}
        except subprocess.TimeoutExpired:
            res = {"ok": False, "timeout": True}
        res["seconds"] = round(time.time() - t0, 1)
        results[name] = res
        print(("PASS " if res.get("ok") else "FAIL ") + name + " " + json.dumps(res)[:700], flush=True)
        with open(os.path.join(ROOT, "gpurun_out", "diag.json"), "w") as f:
            json.dump(results, f, indent=1)
    n_ok = sum(1 for r in results.values() if r.get("ok"))
    print(f"diag: {n_ok}/{len(results)} passed")


if __name__ == "__main__":
    main()
