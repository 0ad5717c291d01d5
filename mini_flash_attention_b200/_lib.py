"""ctypes binding of libmfa_b200.so (the C-ABI in include/mfa_b200.h).

There is no fallback: if the shared library is missing or does not export the ABI, importing the
compute entry points raises.  The product path never touches oracle/.
"""
from __future__ import annotations

import ctypes
import os

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# MFA_B200_LIB: kernel-tuning hook (scripts/tune_prefill.py) to load an alternative build of the same ABI
LIB_PATH = os.environ.get("MFA_B200_LIB") or os.path.join(_PKG_DIR, "libmfa_b200.so")

MFA_OK = 0
ABI_VERSION = 4


class MfaParams(ctypes.Structure):
    """Mirror of mfa_params_t (include/mfa_b200.h), itself a POD mirror of the reference's
    ForwardParams (/root/reference/csrc/mfa/flash.h:8-73)."""

    _fields_ = [
        ("q_ptr", ctypes.c_void_p), ("k_ptr", ctypes.c_void_p), ("v_ptr", ctypes.c_void_p), ("o_ptr", ctypes.c_void_p),
        ("q_batch_stride", ctypes.c_int64), ("q_head_stride", ctypes.c_int64), ("q_row_stride", ctypes.c_int64),
        ("k_batch_stride", ctypes.c_int64), ("k_head_stride", ctypes.c_int64), ("k_row_stride", ctypes.c_int64),
        ("v_batch_stride", ctypes.c_int64), ("v_head_stride", ctypes.c_int64), ("v_row_stride", ctypes.c_int64),
        ("o_batch_stride", ctypes.c_int64), ("o_head_stride", ctypes.c_int64), ("o_row_stride", ctypes.c_int64),
        ("is_causal", ctypes.c_int32), ("window_size_left", ctypes.c_int32), ("window_size_right", ctypes.c_int32),
        ("heads", ctypes.c_int32), ("kv_heads", ctypes.c_int32), ("batch", ctypes.c_int32),
        ("seqlen_q", ctypes.c_int32), ("seqlen_k", ctypes.c_int32), ("head_dim", ctypes.c_int32),
        ("is_bf16", ctypes.c_int32), ("softmax_scale", ctypes.c_float),
        ("cu_seqlens_q", ctypes.c_void_p), ("cu_seqlens_k", ctypes.c_void_p),
        ("total_q", ctypes.c_int32), ("total_k", ctypes.c_int32),
        ("block_table", ctypes.c_void_p),
        ("block_table_batch_stride", ctypes.c_int32), ("max_blocks_per_seq", ctypes.c_int32),
        ("page_block_size", ctypes.c_int32), ("num_blocks", ctypes.c_int32),
        ("k_cache_block_stride", ctypes.c_int64), ("v_cache_block_stride", ctypes.c_int64),
        ("seqlens_k", ctypes.c_void_p),
        ("num_splits", ctypes.c_int32),
        ("softmax_lse_ptr", ctypes.c_void_p),
        ("partial_out", ctypes.c_int32), ("causal_bottom_right", ctypes.c_int32),
        ("local_window", ctypes.c_int32),
        ("kv_fp8", ctypes.c_int32), ("kv_scale_stride", ctypes.c_int32),
        ("k_scale_ptr", ctypes.c_void_p), ("v_scale_ptr", ctypes.c_void_p),
    ]


class MfaAppendParams(ctypes.Structure):
    """Mirror of mfa_append_params_t (include/mfa_b200.h)."""
    _fields_ = [
        ("k_new", ctypes.c_void_p), ("v_new", ctypes.c_void_p),
        ("knew_batch_stride", ctypes.c_int64), ("knew_row_stride", ctypes.c_int64), ("knew_head_stride", ctypes.c_int64),
        ("vnew_batch_stride", ctypes.c_int64), ("vnew_row_stride", ctypes.c_int64), ("vnew_head_stride", ctypes.c_int64),
        ("k_cache", ctypes.c_void_p), ("v_cache", ctypes.c_void_p),
        ("k_batch_stride", ctypes.c_int64), ("k_row_stride", ctypes.c_int64), ("k_head_stride", ctypes.c_int64),
        ("v_batch_stride", ctypes.c_int64), ("v_row_stride", ctypes.c_int64), ("v_head_stride", ctypes.c_int64),
        ("seqlens_k", ctypes.c_void_p), ("seqlens_out", ctypes.c_void_p), ("block_table", ctypes.c_void_p),
        ("block_table_batch_stride", ctypes.c_int32), ("max_blocks_per_seq", ctypes.c_int32),
        ("page_block_size", ctypes.c_int32), ("seqlen_k", ctypes.c_int32),
        ("batch", ctypes.c_int32), ("s_new", ctypes.c_int32), ("kv_heads", ctypes.c_int32), ("head_dim", ctypes.c_int32),
        ("is_bf16", ctypes.c_int32), ("num_blocks", ctypes.c_int32),
    ]


# every symbol include/mfa_b200.h declares (tests/test_abi.py cross-checks this list against the header)
EXPORTS = (
    "mfa_abi_version", "mfa_last_error", "mfa_device_supported", "mfa_prefill_fwd", "mfa_decode_num_splits",
    "mfa_decode_workspace_bytes", "mfa_decode_fwd", "mfa_combine_partials", "mfa_launch_count", "mfa_selftest_umma",
    "mfa_debug_set_trace", "mfa_selftest_umma2", "mfa_kvcache_append", "mfa_decode_static_workspace_limit", "mfa_decode_fwd_static",
    "mfa_xchg_create", "mfa_xchg_export", "mfa_xchg_connect", "mfa_xchg_connect_ptrs", "mfa_xchg_local_ptr",
    "mfa_xchg_status", "mfa_xchg_clear_status", "mfa_xchg_destroy", "mfa_decode_kv_sharded_workspace_bytes", "mfa_decode_fwd_kv_sharded",
)
XCHG_HANDLE_BYTES = 64
XCHG_MAX_WORLD = 8

_lib = None


def load() -> ctypes.CDLL:
    """Load the in-tree shared library; raises (loudly) when it is absent or stale."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"mini_flash_attention_b200: {LIB_PATH} is missing. Build it with "
            "`python -m mini_flash_attention_b200.build` (needs nvcc; there is no CPU fallback).")
    lib = ctypes.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise RuntimeError(f"mini_flash_attention_b200: {LIB_PATH} does not export {name}; rebuild it")
    P = ctypes.POINTER(MfaParams)
    lib.mfa_abi_version.restype = ctypes.c_int
    lib.mfa_last_error.restype = ctypes.c_char_p
    lib.mfa_device_supported.restype = ctypes.c_int
    lib.mfa_prefill_fwd.argtypes = [P, ctypes.c_void_p]
    lib.mfa_prefill_fwd.restype = ctypes.c_int
    lib.mfa_decode_num_splits.argtypes = [P]
    lib.mfa_decode_num_splits.restype = ctypes.c_int
    lib.mfa_decode_workspace_bytes.argtypes = [P, ctypes.c_int]
    lib.mfa_decode_workspace_bytes.restype = ctypes.c_size_t
    lib.mfa_decode_fwd.argtypes = [P, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.mfa_decode_fwd.restype = ctypes.c_int
    lib.mfa_decode_static_workspace_limit.argtypes = []
    lib.mfa_decode_static_workspace_limit.restype = ctypes.c_size_t
    lib.mfa_decode_fwd_static.argtypes = [P, ctypes.c_void_p]
    lib.mfa_decode_fwd_static.restype = ctypes.c_int
    lib.mfa_combine_partials.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.mfa_combine_partials.restype = ctypes.c_int
    lib.mfa_launch_count.restype = ctypes.c_uint64
    lib.mfa_kvcache_append.argtypes = [ctypes.POINTER(MfaAppendParams), ctypes.c_void_p]
    lib.mfa_kvcache_append.restype = ctypes.c_int
    lib.mfa_xchg_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
    lib.mfa_xchg_create.restype = ctypes.c_int
    lib.mfa_xchg_export.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.mfa_xchg_export.restype = ctypes.c_int
    lib.mfa_xchg_connect.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.mfa_xchg_connect.restype = ctypes.c_int
    lib.mfa_xchg_connect_ptrs.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
    lib.mfa_xchg_connect_ptrs.restype = ctypes.c_int
    lib.mfa_xchg_local_ptr.argtypes = [ctypes.c_void_p]
    lib.mfa_xchg_local_ptr.restype = ctypes.c_void_p
    lib.mfa_xchg_status.argtypes = [ctypes.c_void_p]
    lib.mfa_xchg_status.restype = ctypes.c_int
    lib.mfa_xchg_clear_status.argtypes = [ctypes.c_void_p]
    lib.mfa_xchg_clear_status.restype = ctypes.c_int
    lib.mfa_xchg_destroy.argtypes = [ctypes.c_void_p]
    lib.mfa_xchg_destroy.restype = ctypes.c_int
    lib.mfa_decode_kv_sharded_workspace_bytes.argtypes = [P, ctypes.c_int]
    lib.mfa_decode_kv_sharded_workspace_bytes.restype = ctypes.c_size_t
    lib.mfa_decode_fwd_kv_sharded.argtypes = [P, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
    lib.mfa_decode_fwd_kv_sharded.restype = ctypes.c_int
    lib.mfa_debug_set_trace.argtypes = [ctypes.c_void_p]
    lib.mfa_debug_set_trace.restype = None
    lib.mfa_selftest_umma.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]
    lib.mfa_selftest_umma.restype = ctypes.c_int
    lib.mfa_selftest_umma2.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                       ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.mfa_selftest_umma2.restype = ctypes.c_int
    if lib.mfa_abi_version() != ABI_VERSION:
        raise RuntimeError(f"mini_flash_attention_b200: ABI version mismatch ({lib.mfa_abi_version()} != {ABI_VERSION})")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    """Turn a C-ABI error code into the reference's error convention (TORCH_CHECK => RuntimeError)."""
    if rc != MFA_OK:
        msg = load().mfa_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")


def launch_count() -> int:
    return int(load().mfa_launch_count())
