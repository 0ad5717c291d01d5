#!/usr/bin/env python
"""Generate golden input/output vectors from the UNMODIFIED reference CUDA kernels.

Run on a B200 box (the reference has no CPU path):

    python tests/golden/make_golden.py            # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/*.npz tests/golden/      # then commit

It loads oracle/_ref/mini_flash_attention_ref_C.so (built from /root/reference by oracle/build_ref.py, no
source copied) and calls its three pybind functions (/root/reference/csrc/api.cpp:6-8) directly -- the
reference's Python wrapper is bypassed only because it rejects the `causal` argument its own tests pass
(observation 1).  Inputs are seeded on the CPU generator and stored next to the
outputs, so the fixtures do not depend on RNG reproducibility.

Only cases where the reference is known to be correct are generated (page sizes that are multiples of 64,
Sq == Sk for causal, head dims that are powers of two): see SURVEY.md Appendix A, Q1/Q3/Q14.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OUT = os.path.join(ROOT, "gpurun_out", "golden")


def load_ref():
    so = os.path.join(ROOT, "oracle", "_ref", "mini_flash_attention_ref_C.so")
    spec = importlib.util.spec_from_file_location("mini_flash_attention_ref_C", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def to_np(t):
    t = t.detach().cpu()
    if t.dtype == torch.bfloat16:
        return t.float().numpy()
    return t.numpy()


def main():
    ref = load_ref()
    os.makedirs(OUT, exist_ok=True)
    g = torch.Generator().manual_seed(20261104)
    dev = "cuda"

    def rnd(*shape, dtype=torch.float16):
        return torch.randn(*shape, generator=g).to(dtype)

    # ---- fixed-length prefill
    for name, (B, S, H, Hkv, D, causal, dt) in {
        "prefill_s200_gqa_d64": (1, 200, 4, 2, 64, False, torch.float16),
        "prefill_s200_gqa_d64_causal": (1, 200, 4, 2, 64, True, torch.float16),
        "prefill_s130_d128_causal_bf16": (2, 130, 2, 2, 128, True, torch.bfloat16),
        "prefill_s65_d32": (1, 65, 2, 1, 32, False, torch.float16),
        "prefill_s96_d256_causal": (1, 96, 2, 2, 256, True, torch.float16),
    }.items():
        q, k, v = rnd(B, S, H, D, dtype=dt), rnd(B, S, Hkv, D, dtype=dt), rnd(B, S, Hkv, D, dtype=dt)
        o = ref.mini_flash_attention_forward(q.to(dev), k.to(dev), v.to(dev), None, causal, -1, -1)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), kind="prefill", q=to_np(q), k=to_np(k), v=to_np(v),
                            out=to_np(o), causal=causal, bf16=(dt == torch.bfloat16))
        print(name, tuple(o.shape), float(o.float().abs().max()))

    # ---- packed varlen (non-paged and page 64)
    seqlens = [17, 64, 100]
    cu = torch.tensor([0] + list(np.cumsum(seqlens)), dtype=torch.int32)
    total, H, Hkv, D = sum(seqlens), 4, 2, 64
    q, k, v = rnd(total, H, D), rnd(total, Hkv, D), rnd(total, Hkv, D)
    o = ref.mini_flash_attention_varlen_forward(q.to(dev), k.to(dev), v.to(dev), cu.to(dev), cu.to(dev), max(seqlens),
                                                max(seqlens), True, -1, -1, None)
    np.savez_compressed(os.path.join(OUT, "varlen_causal_gqa_d64.npz"), kind="varlen", q=to_np(q), k=to_np(k), v=to_np(v),
                        cu_q=cu.numpy(), cu_k=cu.numpy(), out=to_np(o), causal=True, bf16=False)
    print("varlen", tuple(o.shape))

    page, nblk = 64, 8
    kq = [5, 40, 64]            # query lengths
    kk = [5, 40, 64]            # key lengths (== query lengths: top-left and bottom-right causal agree)
    cu_q = torch.tensor([0] + list(np.cumsum(kq)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(np.cumsum(kk)), dtype=torch.int32)
    q = rnd(sum(kq), H, D)
    kc, vc = rnd(nblk, page, Hkv, D), rnd(nblk, page, Hkv, D)
    bt = torch.tensor([[6, 0], [3, 0], [1, 0]], dtype=torch.int32)
    o = ref.mini_flash_attention_varlen_forward(q.to(dev), kc.to(dev), vc.to(dev), cu_q.to(dev), cu_k.to(dev), max(kq),
                                                max(kk), True, -1, -1, bt.to(dev))
    np.savez_compressed(os.path.join(OUT, "varlen_paged64_causal.npz"), kind="varlen_paged", q=to_np(q), k=to_np(kc),
                        v=to_np(vc), cu_q=cu_q.numpy(), cu_k=cu_k.numpy(), block_table=bt.numpy(), out=to_np(o),
                        causal=True, bf16=False)
    print("varlen paged", tuple(o.shape))

    # ---- decode (paged 256, ragged lengths, 1 and 2 splits)
    B, H, Hkv, D, page, per = 3, 8, 2, 128, 256, 2
    q = rnd(B, 1, H, D)
    nblk = B * per + 1
    kc, vc = rnd(nblk, page, Hkv, D), rnd(nblk, page, Hkv, D)
    bt = torch.tensor([[4, 1], [0, 6], [2, 5]], dtype=torch.int32)
    lens = torch.tensor([300, 257, 64], dtype=torch.int32)
    for splits in (1, 2):
        o = ref.mini_flash_attention_with_kvcache(q.to(dev), kc.to(dev), vc.to(dev), lens.to(dev), bt.to(dev), False, splits)
        np.savez_compressed(os.path.join(OUT, f"decode_paged256_gqa_splits{splits}.npz"), kind="decode", q=to_np(q),
                            k=to_np(kc), v=to_np(vc), block_table=bt.numpy(), seqlens=lens.numpy(), out=to_np(o),
                            splits=splits, bf16=False)
        print("decode", splits, tuple(o.shape))
    torch.cuda.synchronize()
    print("golden written to", OUT)


if __name__ == "__main__":
    sys.exit(main())
