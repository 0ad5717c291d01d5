"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports exactly what
include/mfa_b200.h declares; the Python surface matches the reference's; validation errors follow the
reference's convention (TORCH_CHECK => RuntimeError, /root/reference/csrc/mfa/api.cpp:126-162)."""
import ctypes
import inspect
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "mfa_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mfa_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_header_symbol(mfa_lib):
    from mini_flash_attention_b200 import _lib
    syms = _header_symbols()
    assert len(syms) >= 9
    for s in syms:
        assert hasattr(mfa_lib, s), f"{s} declared in include/mfa_b200.h but not exported"
    assert sorted(_lib.EXPORTS) == syms
    assert mfa_lib.mfa_abi_version() == 4


def test_params_struct_matches_header_layout(mfa_lib):
    """ctypes mirror and the C struct must agree: compile a tiny C program that prints sizeof/offsetof."""
    import subprocess
    import tempfile
    from mini_flash_attention_b200._lib import MfaParams
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "mfa_b200.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu\n", sizeof(mfa_params_t), offsetof(mfa_params_t, is_causal),
         offsetof(mfa_params_t, softmax_scale), offsetof(mfa_params_t, cu_seqlens_q),
         offsetof(mfa_params_t, k_cache_block_stride), offsetof(mfa_params_t, partial_out));
  printf("%zu %zu %zu\n", offsetof(mfa_params_t, kv_fp8), offsetof(mfa_params_t, k_scale_ptr), offsetof(mfa_params_t, v_scale_ptr));
  printf("%zu %zu %zu %zu %zu %zu\n", offsetof(mfa_params_t, local_window), offsetof(mfa_params_t, causal_bottom_right), sizeof(mfa_append_params_t),
         offsetof(mfa_append_params_t, k_cache), offsetof(mfa_append_params_t, seqlens_out),
         offsetof(mfa_append_params_t, is_bf16));
  printf("%zu\n", offsetof(mfa_append_params_t, num_blocks));
  return 0; }
'''
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "t")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        vals = [int(x) for x in subprocess.check_output([exe]).split()]
    from mini_flash_attention_b200._lib import MfaAppendParams as A
    assert vals == [ctypes.sizeof(MfaParams), MfaParams.is_causal.offset, MfaParams.softmax_scale.offset,
                    MfaParams.cu_seqlens_q.offset, MfaParams.k_cache_block_stride.offset, MfaParams.partial_out.offset,
                    MfaParams.kv_fp8.offset, MfaParams.k_scale_ptr.offset, MfaParams.v_scale_ptr.offset,
                    MfaParams.local_window.offset, MfaParams.causal_bottom_right.offset, ctypes.sizeof(A), A.k_cache.offset, A.seqlens_out.offset,
                    A.is_bf16.offset, A.num_blocks.offset]


def test_python_surface_matches_reference():
    """Signatures verbatim from /root/reference/mini_flash_attention/interface.py:6-11,40-50,95-102
    (+ the compatible kw-only `causal` on kvcache, SURVEY section 4 observation 1, and the kw-only, default-off
    `return_lse` of SURVEY 8f row 1): positional parameters are exactly the reference's."""
    def positional(f):
        return [n for n, p in inspect.signature(f).parameters.items() if p.kind is not inspect.Parameter.KEYWORD_ONLY]

    def kw_only(f):
        return {n: p.default for n, p in inspect.signature(f).parameters.items() if p.kind is inspect.Parameter.KEYWORD_ONLY}
    import mini_flash_attention as m
    assert m.__all__ == ["flash_attn_func", "flash_attn_varlen_func", "flash_attn_with_kvcache"]
    assert positional(m.flash_attn_func) == ["q", "k", "v", "causal"]
    assert kw_only(m.flash_attn_func) == {"return_lse": False, "bottom_right": False, "window_size": (-1, -1)}
    assert positional(m.flash_attn_varlen_func) == [
        "q", "k", "v", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k", "causal", "block_table"]
    assert kw_only(m.flash_attn_varlen_func) == {"return_lse": False, "bottom_right": False, "window_size": (-1, -1)}
    sig = inspect.signature(m.flash_attn_with_kvcache)
    assert positional(m.flash_attn_with_kvcache) == ["q", "k_cache", "v_cache", "cache_seqlens", "block_table", "num_splits"]
    assert kw_only(m.flash_attn_with_kvcache) == {"causal": False, "return_lse": False, "k": None, "v": None,
                                                     "window_size": (-1, -1), "k_scale": None, "v_scale": None}
    assert sig.parameters["num_splits"].default == 0
    for name in ("mini_flash_attention_forward", "mini_flash_attention_varlen_forward", "mini_flash_attention_with_kvcache"):
        assert hasattr(m._C, name)


def test_no_cpu_fallback(mfa_lib):
    """The product path refuses CPU tensors like the reference's CHECK_DEVICE; it never computes on the host."""
    import mini_flash_attention as m
    q = torch.randn(1, 8, 2, 64).half()
    with pytest.raises(RuntimeError, match="CUDA"):
        m.flash_attn_func(q, q, q)


def test_validation_errors_are_runtime_errors(mfa_lib):
    import mini_flash_attention as m
    q = torch.randn(1, 8, 2, 64)
    with pytest.raises(RuntimeError, match="fp16 and bf16"):
        m.flash_attn_func(q, q, q)
    with pytest.raises(AssertionError):  # interface.py:116
        m.flash_attn_with_kvcache(torch.randn(1, 2, 2, 64).half(), q.half(), q.half())


def test_product_package_does_not_import_oracle():
    """oracle/ is test infrastructure: nothing under mini_flash_attention_b200/ may reference it."""
    pkg = os.path.join(ROOT, "mini_flash_attention_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"(^|\n)\s*(import|from)\s+oracle|liboracle|attention_oracle|ref_attention", text), \
                    f"{f} uses the oracle"


def test_c_abi_rejects_bad_arguments_without_a_gpu(mfa_lib):
    """Argument validation happens before any CUDA call, so it is checkable on the CPU box."""
    from mini_flash_attention_b200._lib import MfaParams
    P = MfaParams()
    assert mfa_lib.mfa_prefill_fwd(ctypes.byref(P), None) != 0
    assert b"non-NULL" in mfa_lib.mfa_last_error()
    P.q_ptr = P.k_ptr = P.v_ptr = P.o_ptr = 16
    P.head_dim, P.heads, P.kv_heads, P.batch = 512, 8, 2, 1
    assert mfa_lib.mfa_prefill_fwd(ctypes.byref(P), None) == 1
    assert b"256" in mfa_lib.mfa_last_error()
    P.head_dim, P.kv_heads = 64, 3
    assert mfa_lib.mfa_decode_fwd(ctypes.byref(P), None, 0, None) == 1
    assert b"divisible" in mfa_lib.mfa_last_error()
    P.kv_heads, P.seqlen_k, P.num_splits = 2, 8192, 0
    assert mfa_lib.mfa_decode_num_splits(ctypes.byref(P)) >= 1
    P.num_splits = 1000
    assert mfa_lib.mfa_decode_num_splits(ctypes.byref(P)) == 128  # clamp to ceil(8192/64), api.cpp:325-327
    assert mfa_lib.mfa_decode_workspace_bytes(ctypes.byref(P), 4) == 4 * 8 * 65 * 4


@pytest.mark.parametrize("B,H,Hkv,KV,want", [
    (24, 32, 8, 512, 1), (24, 32, 8, 2048, 1), (24, 32, 8, 4096, 1),    # 192 items: unsplit ...
    (24, 32, 8, 8192, 3),                                                # ... except long items cut into two full waves
    (24, 24, 24, 8192, 1),                                               # 576 items: already two waves
    (8, 32, 8, 1024, 2), (8, 32, 8, 4096, 4), (8, 32, 8, 8192, 4),      # one wave: slots / items
    (4, 64, 8, 8192, 9),
    (1, 32, 8, 65536, 37), (1, 32, 8, 1 << 20, 37),                     # long context: 296 slots / 8 KV heads
    (1, 32, 8, 256, 1),                                                  # fewer than 8 tiles per split: no split
])
def test_decode_split_heuristic(mfa_lib, B, H, Hkv, KV, want):
    """The split count fitted on the L2-cold sweep (scripts/decode_splits_gqa.py): splits =
    min(slots / items, tiles / 8) with slots = 2 x 148.  Runs without a GPU (the SM count falls back to a B200's 148)."""
    from mini_flash_attention_b200._lib import MfaParams
    P = MfaParams()
    P.q_ptr = P.k_ptr = P.v_ptr = P.o_ptr = 16
    P.head_dim, P.heads, P.kv_heads, P.batch, P.seqlen_q, P.seqlen_k, P.num_splits = 128, H, Hkv, B, 1, KV, 0
    if torch.cuda.is_available() and torch.cuda.get_device_properties(0).multi_processor_count != 148:
        pytest.skip("fitted for 148 SMs")
    assert mfa_lib.mfa_decode_num_splits(ctypes.byref(P)) == want
