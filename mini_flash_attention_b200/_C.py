"""Host boundary: the three extension functions of the reference, re-implemented over the C-ABI.

Same names, positional signatures, argument meaning and error behaviour as the reference's pybind module
``mini_flash_attention._C`` (/root/reference/csrc/api.cpp:4-9, typed in mini_flash_attention/_C.pyi:4-41),
whose bodies live in /root/reference/csrc/mfa/api.cpp:

    mini_flash_attention_forward          <- mfa::flash_attention_forward          api.cpp:113-186
    mini_flash_attention_varlen_forward   <- mfa::flash_attention_varlen_forward   api.cpp:189-267
    mini_flash_attention_with_kvcache     <- mfa::mha_fwd_kvcache                  api.cpp:343-446

PyTorch is used for device memory (outputs, split workspaces through the caching allocator), the current
stream and the device guard -- plumbing only.  All arithmetic happens in libmfa_b200.so (hand-written
sm_100a kernels).  There is no CPU path: non-CUDA tensors raise RuntimeError like the reference's
CHECK_DEVICE (api.cpp:7, 131-133).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Union

import torch

from . import _lib
from ._lib import MfaParams


def _check(cond: bool, msg: str) -> None:
    if not cond:  # TORCH_CHECK => RuntimeError
        raise RuntimeError(msg)


def _check_qkv(q, k, v) -> None:
    # api.cpp:125-137
    _check(q.dtype in (torch.float16, torch.bfloat16), "FlashAttention only support fp16 and bf16 data type")
    _check(k.dtype == q.dtype, "query and key must have the same dtype")
    _check(v.dtype == q.dtype, "query and value must have the same dtype")
    _check(q.is_cuda, "q must be on CUDA")
    _check(k.is_cuda, "k must be on CUDA")
    _check(v.is_cuda, "v must be on CUDA")
    _check(q.stride(-1) == 1, "Input tensor must have contiguous last dimension")
    _check(k.stride(-1) == 1, "Input tensor must have contiguous last dimension")
    _check(v.stride(-1) == 1, "Input tensor must have contiguous last dimension")


def _tma_ok(t: torch.Tensor) -> bool:
    """TMA needs a 16-byte aligned base and 16-byte multiples for every non-innermost stride."""
    if t.data_ptr() % 16:
        return False
    if t.element_size() == 1:   # e4m3 cache: 16 elements per 16 bytes
        return all(s % 16 == 0 and (s > 0 or n <= 1) for s, n in zip(t.stride()[:-1], t.shape[:-1])) and t.stride(-1) == 1
    # a broadcast view (expand(): stride 0 on an extent > 1) cannot be expressed by a tensor map either
    return all(s % 8 == 0 and (s > 0 or n <= 1) for s, n in zip(t.stride()[:-1], t.shape[:-1])) and t.stride(-1) == 1


def _tma_ready(t: torch.Tensor) -> torch.Tensor:
    return t if _tma_ok(t) else t.contiguous()


def _check_head_dim(head_dim: int, num_heads: int, kv_num_heads: int) -> None:
    _check(head_dim <= 256, "head dimension must be less than or equal to 256")  # api.cpp:152
    _check(kv_num_heads > 0 and num_heads % kv_num_heads == 0,
           "number of key/value heads must be divisible by number of query heads")  # api.cpp:153
    # The reference rounds head_dim up to a multiple of 32 and then reads the neighbouring head
    # (static_switch.h:25-52); we refuse instead of returning wrong numbers.
    _check(head_dim % 8 == 0, "head dimension must be a multiple of 8")


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _i32_ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def mini_flash_attention_forward(q, k, v, out, is_causal, window_size_left, window_size_right, _want_lse=False,
                                 _bottom_right=False, _local_window=False):
    """Fixed-length prefill.  q (B,Sq,H,D); k, v (B,Sk,Hkv,D); returns (B,Sq,H,D).  api.cpp:113-186.
    With _want_lse (the `_lse` entry point) also the natural-log LSE (B,H,Sq) fp32."""
    _check_qkv(q, k, v)
    _check(q.dim() == 4 and k.dim() == 4 and v.dim() == 4, "q, k, v must be 4-D (batch, seqlen, heads, head_dim)")
    batch, seqlen_q, num_heads, head_dim = q.shape
    seqlen_k, kv_num_heads = k.shape[1], k.shape[2]
    _check(k.shape[0] == batch, "batch size of q and k must be the same")
    _check(v.shape[0] == batch, "batch size of q and v must be the same")
    _check(v.shape[1] == seqlen_k, "sequence length of k must be the same as v")
    _check_head_dim(head_dim, num_heads, kv_num_heads)
    _check(tuple(k.shape) == (batch, seqlen_k, kv_num_heads, head_dim), "k must have shape (batch, seqlen_k, kv_num_heads, head_dim)")
    _check(tuple(v.shape) == (batch, seqlen_k, kv_num_heads, head_dim), "v must have shape (batch, seqlen_k, kv_num_heads, head_dim)")
    if out is not None:
        _check(out.stride(-1) == 1, "Output tensor must have contiguous last dimension")
        _check(tuple(out.shape) == (batch, seqlen_q, num_heads, head_dim), "out must have shape (batch, seqlen_q, num_heads, head_dim)")
        _check(out.dtype == q.dtype and out.is_cuda, "out must match q's dtype and device")
    else:
        out = torch.empty_like(q)
    lse = torch.empty(batch, num_heads, seqlen_q, dtype=torch.float32, device=q.device) if _want_lse else None
    if batch == 0 or seqlen_q == 0 or num_heads == 0:
        return (out, lse) if _want_lse else out
    if seqlen_k == 0:
        return (out.zero_(), lse.fill_(float("-inf"))) if _want_lse else out.zero_()
    qc, kc, vc = _tma_ready(q), _tma_ready(k), _tma_ready(v)
    oc = out if _tma_ok(out) else torch.empty(out.shape, dtype=out.dtype, device=out.device)

    P = MfaParams()
    P.q_ptr, P.k_ptr, P.v_ptr, P.o_ptr = qc.data_ptr(), kc.data_ptr(), vc.data_ptr(), oc.data_ptr()
    P.q_batch_stride, P.q_row_stride, P.q_head_stride = qc.stride(0), qc.stride(1), qc.stride(2)
    P.k_batch_stride, P.k_row_stride, P.k_head_stride = kc.stride(0), kc.stride(1), kc.stride(2)
    P.v_batch_stride, P.v_row_stride, P.v_head_stride = vc.stride(0), vc.stride(1), vc.stride(2)
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = oc.stride(0), oc.stride(1), oc.stride(2)
    P.is_causal = int(bool(is_causal))  # api.cpp:157, :88
    P.causal_bottom_right = int(bool(_bottom_right))  # new; 0 = the reference's top-left mask
    P.local_window = int(bool(_local_window))  # new; 0 = window sizes ignored like the reference
    P.window_size_left, P.window_size_right = int(window_size_left), int(window_size_right)
    P.heads, P.kv_heads, P.batch = num_heads, kv_num_heads, batch
    P.seqlen_q, P.seqlen_k, P.head_dim = seqlen_q, seqlen_k, head_dim
    P.is_bf16 = int(q.dtype == torch.bfloat16)
    if _want_lse:
        P.softmax_lse_ptr = lse.data_ptr()
    with torch.cuda.device(q.device):  # api.cpp:123
        _lib.check(_lib.load().mfa_prefill_fwd(ctypes.byref(P), _stream_ptr(q.device)), "mfa_prefill_fwd")
    if oc is not out:
        out.copy_(oc)
    return (out, lse) if _want_lse else out


def mini_flash_attention_varlen_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal,
                                        window_size_left, window_size_right, block_table, _want_lse=False,
                                        _bottom_right=False, _local_window=False):
    """Packed variable-length prefill, optionally with paged K/V.  api.cpp:189-267.

    q (total_q,H,D); k, v (total_k,Hkv,D) or, with block_table (batch,max_blocks) int32,
    (num_blocks,page,Hkv,D).  Superset of the reference: total_k may differ from total_q (quirk Q4)."""
    _check_qkv(q, k, v)
    _check(cu_seqlens_q.is_cuda, "cu_seqlens_q must be on CUDA")
    _check(cu_seqlens_k.is_cuda, "cu_seqlens_k must be on CUDA")
    _check(cu_seqlens_q.dtype == torch.int32 and cu_seqlens_k.dtype == torch.int32, "cu_seqlens must be int32")
    _check(q.dim() == 3, "q must have shape (total_q, num_heads, head_dim)")
    total_q, num_heads, head_dim = q.shape
    batch = cu_seqlens_q.numel() - 1
    _check(batch >= 0 and cu_seqlens_k.numel() == batch + 1, "cu_seqlens_q and cu_seqlens_k must have batch+1 entries")
    kv_num_heads = k.shape[-2]
    _check_head_dim(head_dim, num_heads, kv_num_heads)
    out = torch.empty_like(q)
    # LSE in flash-attn's packed layout (H, total_q); rows past cu_seqlens_q[batch] are not written by the kernel
    lse = torch.full((num_heads, total_q), float("-inf"), dtype=torch.float32, device=q.device) if _want_lse else None
    if batch == 0 or total_q == 0 or num_heads == 0:
        return (out, lse) if _want_lse else out
    cu_q, cu_k = cu_seqlens_q.contiguous(), cu_seqlens_k.contiguous()

    P = MfaParams()
    if block_table is not None:
        _check(k.dim() == 4 and v.dim() == 4, "paged k/v must have shape (num_blocks, page_block_size, kv_heads, head_dim)")
        num_blocks, page = k.shape[0], k.shape[1]
        _check(tuple(k.shape) == (num_blocks, page, kv_num_heads, head_dim), "k must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)")
        _check(tuple(v.shape) == (num_blocks, page, kv_num_heads, head_dim), "v must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)")
        _check(block_table.is_cuda, "block_table must be on CUDA")
        _check(block_table.dtype == torch.int32, "block_table must be int32")
        _check(block_table.dim() == 2 and block_table.shape[0] == batch, "block_table must have the same batch size as q")
        bt = block_table if block_table.stride(1) == 1 else block_table.contiguous()
        kc, vc = _tma_ready(k), _tma_ready(v)
        P.block_table = bt.data_ptr()
        P.block_table_batch_stride = bt.stride(0)
        P.max_blocks_per_seq = bt.shape[1]
        P.page_block_size, P.num_blocks = page, num_blocks
        P.k_cache_block_stride, P.v_cache_block_stride = kc.stride(0), vc.stride(0)
        P.k_row_stride, P.k_head_stride = kc.stride(1), kc.stride(2)
        P.v_row_stride, P.v_head_stride = vc.stride(1), vc.stride(2)
    else:
        _check(k.dim() == 3 and v.dim() == 3, "k, v must have shape (total_k, kv_num_heads, head_dim)")
        total_k = k.shape[0]
        _check(tuple(k.shape) == (total_k, kv_num_heads, head_dim), "k must have shape (total_k, kv_num_heads, head_dim)")
        _check(tuple(v.shape) == (total_k, kv_num_heads, head_dim), "v must have shape (total_k, kv_num_heads, head_dim)")
        if total_k == 0:
            return (out.zero_(), lse) if _want_lse else out.zero_()
        kc, vc = _tma_ready(k), _tma_ready(v)
        P.total_k = total_k
        P.k_row_stride, P.k_head_stride = kc.stride(0), kc.stride(1)
        P.v_row_stride, P.v_head_stride = vc.stride(0), vc.stride(1)
    qc = _tma_ready(q)
    P.q_ptr, P.k_ptr, P.v_ptr, P.o_ptr = qc.data_ptr(), kc.data_ptr(), vc.data_ptr(), out.data_ptr()
    P.q_row_stride, P.q_head_stride = qc.stride(0), qc.stride(1)
    P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1)
    P.is_causal = int(bool(causal))  # api.cpp:231
    P.causal_bottom_right = int(bool(_bottom_right))
    P.local_window = int(bool(_local_window))
    P.window_size_left, P.window_size_right = int(window_size_left), int(window_size_right)
    P.heads, P.kv_heads, P.batch = num_heads, kv_num_heads, batch
    P.seqlen_q, P.seqlen_k, P.head_dim = int(max_seqlen_q), int(max_seqlen_k), head_dim
    P.is_bf16 = int(q.dtype == torch.bfloat16)
    P.cu_seqlens_q, P.cu_seqlens_k = cu_q.data_ptr(), cu_k.data_ptr()
    P.total_q = total_q
    if _want_lse:
        P.softmax_lse_ptr = lse.data_ptr()
    with torch.cuda.device(q.device):
        _lib.check(_lib.load().mfa_prefill_fwd(ctypes.byref(P), _stream_ptr(q.device)), "mfa_prefill_fwd")
    return (out, lse) if _want_lse else out


_FP8 = getattr(torch, "float8_e4m3fn", None)


_scale_cache = {}


def _kv_scale(s, kv_heads: int, device):
    """k_scale / v_scale of an e4m3 cache -> (kv_heads,) fp32 device tensor (None stays None = 1.0).  No kernel launch on
    the per-step path: a Python float maps to a cached device tensor, a (kv_heads,) fp32 tensor is used as it is."""
    if s is None:
        return None
    if not isinstance(s, torch.Tensor):
        key = (float(s), kv_heads, device)
        t = _scale_cache.get(key)
        if t is None:
            if len(_scale_cache) >= 64:
                _scale_cache.clear()
            t = _scale_cache[key] = torch.full((kv_heads,), float(s), dtype=torch.float32, device=device)
        return t
    _check(s.is_cuda, "k_scale / v_scale must be on CUDA (or Python floats)")
    _check(s.numel() in (1, kv_heads), "k_scale / v_scale must have 1 or kv_num_heads elements")
    if s.dtype == torch.float32 and s.numel() == kv_heads and s.is_contiguous():
        return s.reshape(-1)
    return s.to(torch.float32).reshape(-1).expand(kv_heads).contiguous()


def _decode_params(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, k_scale=None, v_scale=None):
    """Validation + parameter block shared by the decode entry points.  api.cpp:343-433."""
    fp8 = _FP8 is not None and k_cache.dtype == _FP8
    if fp8:
        # e4m3 cache (the reference gates the dtype at api.cpp:126-129): q keeps its 16-bit type
        _check(q.dtype in (torch.float16, torch.bfloat16), "FlashAttention only support fp16 and bf16 data type")
        _check(v_cache.dtype == _FP8, "k_cache and v_cache must have the same dtype")
        _check(q.is_cuda and k_cache.is_cuda and v_cache.is_cuda, "q, k_cache and v_cache must be on CUDA")
        _check(q.stride(-1) == 1 and k_cache.stride(-1) == 1 and v_cache.stride(-1) == 1,
               "Input tensor must have contiguous last dimension")
        _check(q.shape[-1] % 128 == 0, "an e4m3 KV cache needs head_dim to be a multiple of 128")
    else:
        _check(k_scale is None and v_scale is None, "k_scale / v_scale are for an e4m3 (torch.float8_e4m3fn) cache")
        _check_qkv(q, k_cache, v_cache)
    _check(q.dim() == 4, "q must have shape (batch, 1, num_heads, head_dim)")
    batch, seqlen_q, num_heads, head_dim = q.shape
    _check(seqlen_q == 1, f"flash decoding expects seqlen_q == 1, got {seqlen_q}")  # api.cpp:377
    kv_num_heads = k_cache.shape[-2]
    _check_head_dim(head_dim, num_heads, kv_num_heads)
    _check(k_cache.dim() == 4 and v_cache.dim() == 4, "k_cache / v_cache must be 4-D")
    keep = []
    P = MfaParams()
    kc, vc = _tma_ready(k_cache), _tma_ready(v_cache)
    if block_table is not None:
        _check(block_table.is_cuda, "block_table must be on CUDA")
        _check(block_table.dtype == torch.int32, "block_table must be int32")
        _check(block_table.dim() == 2 and block_table.shape[0] == batch, "block_table must have the same batch size as q")
        num_blocks, page = k_cache.shape[0], k_cache.shape[1]
        _check(tuple(k_cache.shape) == (num_blocks, page, kv_num_heads, head_dim), "k_cache must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)")
        _check(tuple(v_cache.shape) == (num_blocks, page, kv_num_heads, head_dim), "v_cache must have shape (num_blocks, page_block_size, kv_num_heads, head_dim)")
        bt = block_table if block_table.stride(1) == 1 else block_table.contiguous()
        keep.append(bt)
        seqlen_k = bt.shape[1] * page  # api.cpp:392-394
        P.block_table = bt.data_ptr()
        P.block_table_batch_stride = bt.stride(0)
        P.max_blocks_per_seq = bt.shape[1]
        P.page_block_size, P.num_blocks = page, num_blocks
        P.k_cache_block_stride, P.v_cache_block_stride = kc.stride(0), vc.stride(0)
    else:
        seqlen_k = k_cache.shape[1]
        _check(tuple(k_cache.shape) == (batch, seqlen_k, kv_num_heads, head_dim), "k_cache must have shape (batch, seqlen_k, kv_num_heads, head_dim)")
        _check(tuple(v_cache.shape) == (batch, seqlen_k, kv_num_heads, head_dim), "v_cache must have shape (batch, seqlen_k, kv_num_heads, head_dim)")
        P.k_batch_stride, P.v_batch_stride = kc.stride(0), vc.stride(0)
    P.k_row_stride, P.k_head_stride = kc.stride(1), kc.stride(2)
    P.v_row_stride, P.v_head_stride = vc.stride(1), vc.stride(2)
    # cache_seqlens: tensor (reference), or -- superset, SURVEY quirk Q5 -- None (full cache) / int (broadcast)
    if cache_seqlens is None:
        P.seqlens_k = None
    else:
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((batch,), cache_seqlens, dtype=torch.int32, device=q.device)
        _check(cache_seqlens.is_cuda, "seqlens_k must be on CUDA")
        _check(cache_seqlens.dtype == torch.int32, "seqlens_k must be int32")  # api.cpp:430
        _check(cache_seqlens.numel() == batch, "seqlens_k must have the same number of elements as batch size")
        sl = cache_seqlens.contiguous()
        keep.append(sl)
        P.seqlens_k = sl.data_ptr()
    qc = q if (q.data_ptr() % 4 == 0 and q.stride(0) % 2 == 0 and q.stride(2) % 2 == 0) else q.contiguous()
    keep += [qc, kc, vc]
    P.q_ptr, P.k_ptr, P.v_ptr = qc.data_ptr(), kc.data_ptr(), vc.data_ptr()
    P.q_batch_stride, P.q_row_stride, P.q_head_stride = qc.stride(0), qc.stride(1), qc.stride(2)
    P.heads, P.kv_heads, P.batch = num_heads, kv_num_heads, batch
    P.seqlen_q, P.seqlen_k, P.head_dim = 1, seqlen_k, head_dim
    P.is_bf16 = int(q.dtype == torch.bfloat16)
    P.num_splits = int(num_splits)
    if fp8:
        ks, vs = _kv_scale(k_scale, kv_num_heads, q.device), _kv_scale(v_scale, kv_num_heads, q.device)
        keep += [ks, vs]
        P.kv_fp8, P.kv_scale_stride = 1, 1
        P.k_scale_ptr = ks.data_ptr() if ks is not None else None
        P.v_scale_ptr = vs.data_ptr() if vs is not None else None
    return P, keep


def mini_flash_attention_with_kvcache(q, k_cache, v_cache, cache_seqlens: Optional[Union[int, torch.Tensor]],
                                      block_table, causal, num_splits=0, _want_lse=False, _k_new=None, _v_new=None,
                                      _window_left=-1, _k_scale=None, _v_scale=None):
    """q_len == 1 decode over a contiguous or paged KV cache.  api.cpp:343-446.
    `causal` is accepted and ignored exactly like the reference (api.cpp:349, quirk Q10)."""
    P, keep = _decode_params(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, _k_scale, _v_scale)
    if _window_left is not None and _window_left >= 0:  # sliding window: the last window_left + 1 keys
        P.local_window, P.window_size_left, P.window_size_right = 1, int(_window_left), -1
    if _k_new is not None or _v_new is not None:
        _check(not P.kv_fp8, "appending k/v into an e4m3 cache is not supported (quantise and write the rows yourself)")
        keep.append(_append_kv(P, keep, q, k_cache, v_cache, cache_seqlens, _k_new, _v_new))
    out = torch.empty_like(q)  # api.cpp:405
    # (B, H, 1) fp32: the tensor the reference allocates and drops (api.cpp:436-437)
    lse = torch.empty(q.shape[0], q.shape[2], 1, dtype=torch.float32, device=q.device) if _want_lse else None
    if q.numel() == 0:
        return (out, lse) if _want_lse else out
    if P.seqlen_k == 0:
        return (out.zero_(), lse.fill_(float("-inf"))) if _want_lse else out.zero_()
    oc = out if (out.data_ptr() % 8 == 0 and out.stride(0) % 4 == 0 and out.stride(2) % 4 == 0) else \
        torch.empty(out.shape, dtype=out.dtype, device=out.device)
    P.o_ptr = oc.data_ptr()
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = oc.stride(0), oc.stride(1), oc.stride(2)
    if _want_lse:
        P.softmax_lse_ptr = lse.data_ptr()
    lib = _lib.load()
    with torch.cuda.device(q.device):
        splits = lib.mfa_decode_num_splits(ctypes.byref(P))
        _check(splits >= 1, "mfa_decode_num_splits failed")
        P.num_splits = splits
        ws_bytes = lib.mfa_decode_workspace_bytes(ctypes.byref(P), splits)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=q.device) if ws_bytes else None  # api.cpp:333-334
        _lib.check(lib.mfa_decode_fwd(ctypes.byref(P), ws.data_ptr() if ws is not None else None, ws_bytes,
                                      _stream_ptr(q.device)), "mfa_decode_fwd")
    if oc is not out:
        out.copy_(oc)
    return (out, lse) if _want_lse else out


def _append_kv(P, keep, q, k_cache, v_cache, cache_seqlens, k_new, v_new):
    """KV-cache append: write the new rows in place at cache_seqlens and point the decode
    parameter block at cache_seqlens + s_new (a fresh tensor; cache_seqlens itself is not modified)."""
    _check(k_new is not None and v_new is not None, "k and v must be given together")
    _check(cache_seqlens is not None, "appending k/v needs cache_seqlens (where to write)")
    kc, vc = keep[-2], keep[-1]
    _check(kc is k_cache and vc is v_cache, "appending k/v needs a cache that can be written in place (16-byte aligned, strides multiples of 8)")
    _check(k_new.is_cuda and v_new.is_cuda, "k and v must be on CUDA")
    _check(k_new.dtype == q.dtype and v_new.dtype == q.dtype, "k and v must have q's dtype")
    _check(k_new.dim() == 4 and k_new.shape[0] == q.shape[0] and k_new.shape[2] == P.kv_heads and k_new.shape[3] == P.head_dim,
           "k must have shape (batch, seqlen_new, kv_num_heads, head_dim)")
    _check(v_new.shape == k_new.shape, "v must have the same shape as k")
    kn = _tma_ready(k_new if k_new.stride(-1) == 1 else k_new.contiguous())
    vn = _tma_ready(v_new if v_new.stride(-1) == 1 else v_new.contiguous())
    seqlens_new = torch.empty(q.shape[0], dtype=torch.int32, device=q.device)
    A = _lib.MfaAppendParams()
    A.k_new, A.v_new = kn.data_ptr(), vn.data_ptr()
    A.knew_batch_stride, A.knew_row_stride, A.knew_head_stride = kn.stride(0), kn.stride(1), kn.stride(2)
    A.vnew_batch_stride, A.vnew_row_stride, A.vnew_head_stride = vn.stride(0), vn.stride(1), vn.stride(2)
    A.k_cache, A.v_cache = kc.data_ptr(), vc.data_ptr()
    A.k_batch_stride, A.k_row_stride, A.k_head_stride = kc.stride(0), kc.stride(1), kc.stride(2)
    A.v_batch_stride, A.v_row_stride, A.v_head_stride = vc.stride(0), vc.stride(1), vc.stride(2)
    A.seqlens_k, A.seqlens_out = P.seqlens_k, seqlens_new.data_ptr()
    A.block_table, A.block_table_batch_stride = P.block_table, P.block_table_batch_stride
    A.max_blocks_per_seq, A.page_block_size, A.seqlen_k = P.max_blocks_per_seq, P.page_block_size, P.seqlen_k
    A.batch, A.s_new, A.kv_heads, A.head_dim, A.is_bf16 = P.batch, k_new.shape[1], P.kv_heads, P.head_dim, P.is_bf16
    A.num_blocks = P.num_blocks
    with torch.cuda.device(q.device):
        _lib.check(_lib.load().mfa_kvcache_append(ctypes.byref(A), _stream_ptr(q.device)), "mfa_kvcache_append")
    P.seqlens_k = seqlens_new.data_ptr()
    return (seqlens_new, kn, vn)


def _empty_shard_partial(q, k_cache, block_table):
    """A KV shard with no rows (a rank beyond the end of a short context): O = 0, LSE = -inf, no kernel launch.  The merge
    (merge.cuh) gives such a partial weight exp2(-inf) = 0."""
    if q.numel() == 0 or (block_table is None and k_cache.shape[1] == 0) or k_cache.numel() == 0:
        batch, _, num_heads, head_dim = q.shape
        return (torch.zeros(batch, num_heads, head_dim, dtype=torch.float32, device=q.device),
                torch.full((batch, num_heads), float("-inf"), dtype=torch.float32, device=q.device))
    return None




def decode_partial(q, k_cache, v_cache, cache_seqlens, block_table, num_splits=0):
    """KV-shard building block of the multi-GPU decode: normalised fp32 partial O (B,H,D) and LSE (B,H)
    over this rank's part of the KV axis (LSE = -inf, O = 0 where the shard is empty).  New -- the reference
    has no multi-GPU path; the arithmetic is its split-KV partial (decode.cuh:645-661)."""
    empty = _empty_shard_partial(q, k_cache, block_table)
    if empty is not None:
        return empty
    P, keep = _decode_params(q, k_cache, v_cache, cache_seqlens, block_table, num_splits)
    batch, _, num_heads, head_dim = q.shape
    o = torch.empty(batch, num_heads, head_dim, dtype=torch.float32, device=q.device)
    lse = torch.empty(batch, num_heads, dtype=torch.float32, device=q.device)
    P.o_ptr, P.softmax_lse_ptr, P.partial_out = o.data_ptr(), lse.data_ptr(), 1
    lib = _lib.load()
    with torch.cuda.device(q.device):
        splits = lib.mfa_decode_num_splits(ctypes.byref(P))
        _check(splits >= 1, "mfa_decode_num_splits failed")
        P.num_splits = splits
        ws_bytes = lib.mfa_decode_workspace_bytes(ctypes.byref(P), splits)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=q.device) if ws_bytes else None
        _lib.check(lib.mfa_decode_fwd(ctypes.byref(P), ws.data_ptr() if ws is not None else None, ws_bytes,
                                      _stream_ptr(q.device)), "mfa_decode_fwd")
    return o, lse


def decode_kv_sharded_fused(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, xchg_handle: int,
                            return_lse: bool = False):
    """KV-sharded decode with the exchange fused into the merge kernel (include/mfa_b200.h, csrc/xchg_sm100.cu): this
    rank's rows of the KV axis in, the FULL result (B,1,H,D) in q's dtype out on every rank -- split-KV kernel, then one
    kernel that merges the local splits, pushes the row to the peers over NVLink, waits for theirs and merges.  No NCCL
    call.  `xchg_handle` is a connected mfa_xchg_t* (parallel.KVShardExchange).  New; no reference analogue."""
    P, keep = _decode_params(q, k_cache, v_cache, cache_seqlens, block_table, num_splits)
    batch, _, num_heads, head_dim = q.shape
    _check(P.seqlen_k > 0 and q.numel() > 0, "KV-sharded decode needs a non-empty problem")
    out = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    lse = torch.empty(batch, num_heads, dtype=torch.float32, device=q.device) if return_lse else None
    P.o_ptr = out.data_ptr()
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    if return_lse:
        P.softmax_lse_ptr = lse.data_ptr()
    lib = _lib.load()
    with torch.cuda.device(q.device):
        splits = lib.mfa_decode_num_splits(ctypes.byref(P))
        _check(splits >= 1, "mfa_decode_num_splits failed")
        P.num_splits = splits
        ws_bytes = lib.mfa_decode_kv_sharded_workspace_bytes(ctypes.byref(P), splits)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=q.device)
        _lib.check(lib.mfa_decode_fwd_kv_sharded(ctypes.byref(P), ws.data_ptr(), ws_bytes, xchg_handle,
                                                 _stream_ptr(q.device)), "mfa_decode_fwd_kv_sharded")
    return (out, lse) if return_lse else out


def combine_partials(packed: torch.Tensor, out_dtype: torch.dtype, return_lse: bool = False):
    """LSE merge of R packed partials (R, rows, D+1) fp32 -> (rows, D) out_dtype.  decode.cuh:665-755."""
    _check(packed.is_cuda and packed.dtype == torch.float32 and packed.dim() == 3, "packed must be a CUDA fp32 (R, rows, D+1) tensor")
    _check(out_dtype in (torch.float16, torch.bfloat16), "out dtype must be fp16 or bf16")
    packed = packed.contiguous()
    R, rows, d1 = packed.shape
    out = torch.empty(rows, d1 - 1, dtype=out_dtype, device=packed.device)
    lse = torch.empty(rows, dtype=torch.float32, device=packed.device) if return_lse else None
    with torch.cuda.device(packed.device):
        _lib.check(_lib.load().mfa_combine_partials(packed.data_ptr(), R, rows, d1 - 1, out.data_ptr(), out.stride(0),
                                                    int(out_dtype == torch.bfloat16),
                                                    lse.data_ptr() if lse is not None else None,
                                                    _stream_ptr(packed.device)), "mfa_combine_partials")
    return (out, lse) if return_lse else out


# ---------------------------------------------------------------------------------------------------------------
# Compiled host boundary.  csrc/torch_shim.cpp holds the same validation / allocation / parameter fill in C++ (a torch
# extension module linked against libmfa_b200.so, built in-tree by build.build_shim()); when it is present the three
# entry points and decode_partial dispatch to it, which takes the per-call host cost from ~29 us (ctypes struct fill
# in Python) to a few us.  The Python bodies above stay as the same logic over the same C-ABI: used when the module
# has not been built, when MFA_B200_SHIM=py, or when MFA_B200_LIB points at another build of the library.
# Either way the arithmetic is libmfa_b200.so's CUDA kernels; neither path has a CPU fallback.
# ---------------------------------------------------------------------------------------------------------------
import os as _os

_py_forward = mini_flash_attention_forward
_py_varlen_forward = mini_flash_attention_varlen_forward
_py_with_kvcache = mini_flash_attention_with_kvcache
_py_decode_partial = decode_partial


# (out, lse) variants -- SURVEY 8f row 1: the LSE the reference computes and drops, exposed
def mini_flash_attention_forward_lse(q, k, v, out, is_causal, window_size_left, window_size_right):
    return _py_forward(q, k, v, out, is_causal, window_size_left, window_size_right, _want_lse=True)


def mini_flash_attention_varlen_forward_lse(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal,
                                            window_size_left, window_size_right, block_table):
    return _py_varlen_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal,
                              window_size_left, window_size_right, block_table, _want_lse=True)


def mini_flash_attention_with_kvcache_lse(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits=0):
    return _py_with_kvcache(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits, _want_lse=True)


# superset entry points: optional LSE; mask aligned bottom-right; window sizes HONOURED (sliding window); in-place k/v
# append before the decode
def mini_flash_attention_forward_ext(q, k, v, out, is_causal, window_size_left, window_size_right, return_lse, bottom_right):
    r = _py_forward(q, k, v, out, is_causal, window_size_left, window_size_right, _want_lse=return_lse, _bottom_right=bottom_right,
                    _local_window=True)
    return r if return_lse else (r, None)


def mini_flash_attention_varlen_forward_ext(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal,
                                            window_size_left, window_size_right, block_table, return_lse, bottom_right):
    r = _py_varlen_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal, window_size_left,
                           window_size_right, block_table, _want_lse=return_lse, _bottom_right=bottom_right,
                           _local_window=True)
    return r if return_lse else (r, None)


def mini_flash_attention_with_kvcache_ext(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits, return_lse,
                                          k_new, v_new, window_left=-1):
    r = _py_with_kvcache(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits, _want_lse=return_lse,
                         _k_new=k_new, _v_new=v_new, _window_left=window_left)
    return r if return_lse else (r, None)


def mini_flash_attention_with_kvcache_fp8(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, return_lse,
                                          window_left, k_scale, v_scale):
    """Decode over an e4m3 (torch.float8_e4m3fn) KV cache with per-tensor or per-KV-head dequantisation scales."""
    r = _py_with_kvcache(q, k_cache, v_cache, cache_seqlens, block_table, False, num_splits, _want_lse=return_lse,
                         _window_left=window_left, _k_scale=k_scale, _v_scale=v_scale)
    return r if return_lse else (r, None)


_py_with_kvcache_fp8 = mini_flash_attention_with_kvcache_fp8
_py_forward_ext = mini_flash_attention_forward_ext
_py_varlen_forward_ext = mini_flash_attention_varlen_forward_ext
_py_with_kvcache_ext = mini_flash_attention_with_kvcache_ext
_py_forward_lse = mini_flash_attention_forward_lse
_py_varlen_forward_lse = mini_flash_attention_varlen_forward_lse
_py_with_kvcache_lse = mini_flash_attention_with_kvcache_lse
SHIM = "python"
_shim_error = None


def _load_compiled_shim():
    if _os.environ.get("MFA_B200_SHIM", "").lower() in ("py", "python") or _os.environ.get("MFA_B200_LIB"):
        return None
    global _shim_error
    try:
        from . import _mfa_torch_shim as m  # noqa: WPS433
    except ImportError as e:
        _shim_error = e
        if _os.environ.get("MFA_B200_SHIM", "").lower() in ("cpp", "compiled"):
            raise
        return None
    if not hasattr(m, "mini_flash_attention_with_kvcache_fp8"):   # built from an older torch_shim.cpp: rebuild with build()
        _shim_error = ImportError("stale _mfa_torch_shim module (run python -m mini_flash_attention_b200.build)")
        return None
    return m


_shim = _load_compiled_shim()
if _shim is not None:
    SHIM = "compiled"
    mini_flash_attention_forward = _shim.mini_flash_attention_forward
    mini_flash_attention_varlen_forward = _shim.mini_flash_attention_varlen_forward

    def mini_flash_attention_with_kvcache(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits=0):
        if isinstance(cache_seqlens, int):  # superset (quirk Q5): broadcast an int
            cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=q.device)
        return _shim.mini_flash_attention_with_kvcache(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits)

    def decode_partial(q, k_cache, v_cache, cache_seqlens, block_table, num_splits=0):
        empty = _empty_shard_partial(q, k_cache, block_table)
        if empty is not None:
            return empty
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=q.device)
        return _shim.decode_partial(q, k_cache, v_cache, cache_seqlens, block_table, num_splits)

    mini_flash_attention_forward_lse = _shim.mini_flash_attention_forward_lse
    mini_flash_attention_varlen_forward_lse = _shim.mini_flash_attention_varlen_forward_lse

    def mini_flash_attention_with_kvcache_lse(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits=0):
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=q.device)
        return _shim.mini_flash_attention_with_kvcache_lse(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits)

    mini_flash_attention_forward_ext = _shim.mini_flash_attention_forward_ext
    mini_flash_attention_varlen_forward_ext = _shim.mini_flash_attention_varlen_forward_ext

    def mini_flash_attention_with_kvcache_ext(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits,
                                              return_lse, k_new, v_new, window_left=-1):
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=q.device)
        return _shim.mini_flash_attention_with_kvcache_ext(q, k_cache, v_cache, cache_seqlens, block_table, causal,
                                                           num_splits, return_lse, k_new, v_new, window_left)

    def mini_flash_attention_with_kvcache_fp8(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, return_lse,  # noqa: F811
                                              window_left, k_scale, v_scale):
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=q.device)
        kv_heads = k_cache.shape[-2]
        return _shim.mini_flash_attention_with_kvcache_fp8(q, k_cache, v_cache, cache_seqlens, block_table, num_splits,
                                                           return_lse, window_left, _kv_scale(k_scale, kv_heads, q.device),
                                                           _kv_scale(v_scale, kv_heads, q.device))

    _py_decode_kv_sharded_fused = decode_kv_sharded_fused

    def decode_kv_sharded_fused(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, xchg_handle, return_lse=False):  # noqa: F811
        if return_lse or not hasattr(_shim, "decode_kv_sharded_fused"):
            return _py_decode_kv_sharded_fused(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, xchg_handle, return_lse)
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((q.shape[0],), cache_seqlens, dtype=torch.int32, device=q.device)
        return _shim.decode_kv_sharded_fused(q, k_cache, v_cache, cache_seqlens, block_table, num_splits, xchg_handle)

    mini_flash_attention_with_kvcache.__doc__ = _py_with_kvcache.__doc__
    decode_partial.__doc__ = _py_decode_partial.__doc__
