"""Public Python API -- signatures verbatim from the reference
(/root/reference/mini_flash_attention/interface.py:6-11, :40-50, :95-102).

Each function is, like in the reference, a one-line positional forward into the extension surface
(here ``mini_flash_attention_b200._C``, a C-ABI shim instead of a pybind module).
"""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch

from . import _C


def _window(window_size) -> Tuple[int, int]:
    wl, wr = window_size
    return (int(wl) if wl is not None and wl >= 0 else -1), (int(wr) if wr is not None and wr >= 0 else -1)


def flash_attn_func(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    causal: bool = False,
    *,
    return_lse: bool = False,
    bottom_right: bool = False,
    window_size: Tuple[int, int] = (-1, -1),
):
    """FlashAttention forward (reference: interface.py:6-37).

    Supports MQA/GQA: K, V may have fewer heads than Q (the number of Q heads must be divisible by the number
    of KV heads); Q head h attends to KV head ``h // (Hq // Hkv)``.

    Arguments:
        q: (batch_size, seqlen_q, nheads, headdim), fp16 or bf16, CUDA
        k: (batch_size, seqlen_k, nheads_k, headdim)
        v: (batch_size, seqlen_k, nheads_k, headdim)
        causal: apply the causal mask.  Like the reference it is aligned to the TOP-LEFT corner
            (key j is visible to query i iff j <= i), which equals the usual convention when seqlen_q == seqlen_k.
        bottom_right: (keyword-only, new -- SURVEY 8f row 2, quirk Q1) with causal=True align the mask to the
            BOTTOM-RIGHT corner instead: key j visible to query i iff j <= i + (seqlen_k - seqlen_q), the flash-attn
            >= 2.1 convention and what a chunk of new queries against a longer cache needs.  Default off = reference.
        return_lse: (keyword-only, new -- SURVEY 8f row 1) also return the row-wise log-sum-exp of the scaled scores,
            which the reference computes and drops; makes the op composable (chunked / ring prefill merges).
        window_size: (keyword-only, new -- SURVEY 8f row 3; the reference plumbs window sizes but never implements
            them, flash.h:33-34, README.md:153) sliding-window attention ``(left, right)``: key j is visible to query i
            iff ``i + s - left <= j <= i + s + right`` with s = 0 (top-left, the reference's alignment) or
            ``seqlen_k - seqlen_q`` with bottom_right=True (flash-attn's convention); -1 = unbounded; causal=True
            forces right = 0.  KV tiles outside the band are skipped, so cost scales with the window, not seqlen_k.
    Return:
        out: (batch_size, seqlen_q, nheads, headdim)
        softmax_lse [if return_lse]: (batch_size, nheads, seqlen_q) fp32, natural log; -inf for rows that see no key
    """
    wl, wr = _window(window_size)
    if bottom_right or wl >= 0 or wr >= 0:
        out, lse = _C.mini_flash_attention_forward_ext(q, k, v, None, causal, wl, wr, return_lse, bottom_right)
        return (out, lse) if return_lse else out
    if return_lse:
        return _C.mini_flash_attention_forward_lse(q, k, v, None, causal, -1, -1)
    return _C.mini_flash_attention_forward(q, k, v, None, causal, -1, -1)


def flash_attn_varlen_func(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    cu_seqlens_q: torch.Tensor,
    cu_seqlens_k: torch.Tensor,
    max_seqlen_q: int,
    max_seqlen_k: int,
    causal: bool = False,
    block_table=None,
    *,
    return_lse: bool = False,
    bottom_right: bool = False,
    window_size: Tuple[int, int] = (-1, -1),
):
    """FlashAttention forward for packed variable-length batches (reference: interface.py:40-92).

    Arguments:
        q: (total_q, nheads, headdim)
        k, v: (total_k, nheads_k, headdim), or with ``block_table`` the paged cache
            (num_blocks, page_block_size, nheads_k, headdim)
        cu_seqlens_q, cu_seqlens_k: (batch_size + 1,) int32 cumulative sequence lengths
        max_seqlen_q, max_seqlen_k: maximum sequence lengths in the batch
        causal: top-left aligned causal mask (see flash_attn_func)
        block_table: optional (batch_size, max_blocks_per_seq) int32 page table; pages are gathered per page
            (any page size that is a multiple of 8), not per 64-row tile.
        return_lse: (keyword-only, new) also return the log-sum-exp, see flash_attn_func
        bottom_right: (keyword-only, new) bottom-right aligned causal mask per sequence, see flash_attn_func
        window_size: (keyword-only, new) sliding window (left, right) per sequence, see flash_attn_func
    Return:
        out: (total_q, nheads, headdim)
        softmax_lse [if return_lse]: (nheads, total_q) fp32
    """
    wl, wr = _window(window_size)
    if bottom_right or wl >= 0 or wr >= 0:
        out, lse = _C.mini_flash_attention_varlen_forward_ext(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q,
                                                              max_seqlen_k, causal, wl, wr, block_table, return_lse,
                                                              bottom_right)
        return (out, lse) if return_lse else out
    fn = _C.mini_flash_attention_varlen_forward_lse if return_lse else _C.mini_flash_attention_varlen_forward
    return fn(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, causal, -1, -1, block_table)


def flash_attn_with_kvcache(
    q,
    k_cache,
    v_cache,
    cache_seqlens: Optional[Union[int, torch.Tensor]] = None,
    block_table: Optional[torch.Tensor] = None,
    num_splits=0,
    *,
    causal: bool = False,
    return_lse: bool = False,
    k: Optional[torch.Tensor] = None,
    v: Optional[torch.Tensor] = None,
    window_size: Tuple[int, int] = (-1, -1),
    k_scale: Optional[Union[float, torch.Tensor]] = None,
    v_scale: Optional[Union[float, torch.Tensor]] = None,
):
    """Single-token decode against a KV cache (reference: interface.py:95-124).

    Arguments:
        q: (batch_size, 1, nheads, headdim)
        k_cache, v_cache: (batch_size, seqlen_cache, nheads_k, headdim), or with ``block_table`` the paged cache
            (num_blocks, page_block_size, nheads_k, headdim)
        cache_seqlens: (batch_size,) int32 current lengths; an int is broadcast; None means the full cache
        block_table: optional (batch_size, max_blocks_per_seq) int32
        num_splits: KV splits (flash decoding); <= 0 picks automatically
        causal: accepted for compatibility with the reference's tests and ignored (q_len == 1)
        k, v: (new -- SURVEY 8f row 2; the semantics the reference's docstring describes at interface.py:110-111 but
            does not implement) (batch_size, seqlen_new, nheads_k, headdim) new keys/values: written IN PLACE into the
            cache at positions cache_seqlens[b] .. + seqlen_new - 1 (through block_table when paged) before the
            attention, which then covers cache_seqlens + seqlen_new keys.  cache_seqlens itself is not modified.
            Rows that would fall beyond the cache capacity (or onto a block_table entry outside the cache) are NOT
            written and the attended length is clamped to the capacity: size the cache for the longest sequence.
        window_size: (new -- SURVEY 8f row 3) sliding window (left, right): the query sits at the last position and
            attends to the last ``left + 1`` keys only (right is irrelevant for one query); KV tiles in front of the
            window are not read, so the step costs O(window), not O(cache_seqlens).  -1 = unbounded.
        return_lse: (new) also return softmax_lse (batch_size, nheads, 1) fp32 -- the tensor the reference allocates
            and drops (api.cpp:436-437)
        k_scale, v_scale: (new -- SURVEY 8f row 4) for a cache stored as torch.float8_e4m3fn (half the bytes per key; q
            and the output stay fp16/bf16): dequantisation scales, a float / 1-element tensor (per tensor) or a
            (nheads_k,) fp32 tensor (per KV head); the key is k_cache * k_scale, the value v_cache * v_scale.
            None = 1.  head_dim must be a multiple of 128.  Not combinable with k= / v= (append).
    """
    assert q.size(1) == 1, "flash_attn_with_kvcache currently only supports seqlen_q=1 for decoding"
    wl, _ = _window(window_size)
    if k_cache.dtype not in (torch.float16, torch.bfloat16):   # e4m3 cache
        assert k is None and v is None, "appending k/v into an e4m3 cache is not supported"
        out, lse = _C.mini_flash_attention_with_kvcache_fp8(q, k_cache, v_cache, cache_seqlens, block_table, num_splits,
                                                            return_lse, wl, k_scale, v_scale)
        return (out, lse) if return_lse else out
    assert k_scale is None and v_scale is None, "k_scale / v_scale are for an e4m3 (torch.float8_e4m3fn) cache"
    if k is not None or v is not None or wl >= 0:
        out, lse = _C.mini_flash_attention_with_kvcache_ext(q, k_cache, v_cache, cache_seqlens, block_table, causal,
                                                            num_splits, return_lse, k, v, wl)
        return (out, lse) if return_lse else out
    if return_lse:
        return _C.mini_flash_attention_with_kvcache_lse(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits)
    return _C.mini_flash_attention_with_kvcache(q, k_cache, v_cache, cache_seqlens, block_table, causal, num_splits)
