#!/usr/bin/env python
"""Tuning aid (GPU box): cycles per tcgen05.mma (M=128 per CTA, N=128, K=16, bf16) for SS / TS, single CTA vs CTA pair."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from mini_flash_attention_b200 import _lib
lib = _lib.load()
names = {0: "pair SS", 1: "pair TS(V)", 2: "single SS", 3: "single TS(V)", 4: "pair TS(K)", 5: "single TS(K)"}
for mode in (2, 3, 5, 0, 1, 4):
    a = (torch.randn(256, 128, device="cuda") * 0.01).to(torch.bfloat16)
    b = (torch.randn(128, 128, device="cuda") * 0.01).to(torch.bfloat16)
    d = torch.zeros(256, 128, device="cuda")
    cyc = torch.zeros(2, dtype=torch.int64, device="cuda")
    for reps in (1, 2, 4, 8, 64, 256):
        rc = lib.mfa_selftest_umma2(mode, a.data_ptr(), b.data_ptr(), d.data_ptr(), 1, reps, cyc.data_ptr(), None)
        torch.cuda.synchronize()
        assert rc == 0
        print(f"{names[mode]:13s} reps={reps:4d} cycles={int(cyc[0]):8d} per-MMA={int(cyc[0]) / (reps * 8):7.1f}   issued after {int(cyc[1]):8d} cycles")
