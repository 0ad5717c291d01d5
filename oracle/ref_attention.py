"""CPU oracles for the attention-forward hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference legs) may
import this module; the product package mini_flash_attention_b200 never does.

Two oracles:

* ``attention_ref`` / ``varlen_ref`` / ``kvcache_ref`` -- fp32 PyTorch restatement of the oracle the
  reference's own tests use: softmax(mask(Q K^T / sqrt(D))) V on inputs up-cast from the same
  fp16/bf16 tensors (/root/reference/tests/test_mha.py:75-81, tests/test_causal.py:87-93, GQA by
  repeat_interleave tests/test_gqa.py:118-128), with the REFERENCE's mask rule: causal is
  top-left aligned, key col masked iff col > row (/root/reference/csrc/mfa/prefill.cuh:416).
* ``c_prefill`` / ``c_decode`` -- ctypes front-end of oracle/attention_oracle.c, the tile-by-tile
  C restatement of the reference kernels (online softmax in the exp2 domain, P rounded to the input
  dtype before PV, split-KV + LSE combine).
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from typing import Optional, Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")
_SO = os.path.join(_BUILD, "liboracle.so")


# --------------------------------------------------------------------------- torch fp32 oracle
def _expand_kv(x: torch.Tensor, group: int, dim: int) -> torch.Tensor:
    return x if group == 1 else x.repeat_interleave(group, dim=dim)


def _band_mask(s, Sq, Sk, causal, bottom_right, window_size):
    """-inf outside row + shift - left <= col <= row + shift + right (a bound < 0 absent; causal => right = 0);
    shift = 0 (top-left, the reference's prefill.cuh:416) or Sk - Sq (bottom_right, flash-attn's convention)."""
    wl, wr = window_size
    if causal:
        wr = 0
    if wl < 0 and wr < 0:
        return s
    row = torch.arange(Sq, device=s.device)[:, None] + ((Sk - Sq) if bottom_right else 0)
    col = torch.arange(Sk, device=s.device)[None, :]
    if wr >= 0:
        s = s.masked_fill(col > row + wr, float("-inf"))
    if wl >= 0:
        s = s.masked_fill(col < row - wl, float("-inf"))
    return s


def attention_ref(q, k, v, causal: bool = False, key_len: Optional[int] = None, bottom_right: bool = False,
                  window_size=(-1, -1)) -> torch.Tensor:
    """q (B,Sq,H,D), k/v (B,Sk,Hkv,D) -> (B,Sq,H,D) float32.  Top-left causal (col > row masked, the reference's
    prefill.cuh:416) or, with bottom_right, col > row + (Sk - Sq) masked (flash-attn >= 2.1).
    window_size: sliding-window band (no reference implementation: fp32 math only, parity unpinned)."""
    B, Sq, H, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    qf = q.float().permute(0, 2, 1, 3)
    kf = _expand_kv(k.float(), H // Hkv, 2).permute(0, 2, 1, 3)
    vf = _expand_kv(v.float(), H // Hkv, 2).permute(0, 2, 1, 3)
    s = torch.matmul(qf, kf.transpose(-1, -2)) * (1.0 / math.sqrt(D))
    if key_len is not None and key_len < Sk:
        s[..., key_len:] = float("-inf")
    s = _band_mask(s, Sq, Sk, causal, bottom_right, window_size)
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)  # fully masked rows => 0 output (prefill.cuh:604-605)
    o = torch.matmul(p, vf)
    return o.permute(0, 2, 1, 3).contiguous()


def lse_ref(q, k, causal: bool = False, bottom_right: bool = False, window_size=(-1, -1)) -> torch.Tensor:
    """Row-wise natural-log LSE of the scaled, masked scores: q (B,Sq,H,D), k (B,Sk,Hkv,D) -> (B,H,Sq) float32;
    -inf where a row sees no key.  (The quantity the reference stores in softmax_lse, decode.cuh:750-754.)"""
    B, Sq, H, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    qf = q.float().permute(0, 2, 1, 3)
    kf = _expand_kv(k.float(), H // Hkv, 2).permute(0, 2, 1, 3)
    s = torch.matmul(qf, kf.transpose(-1, -2)) * (1.0 / math.sqrt(D))
    s = _band_mask(s, Sq, Sk, causal, bottom_right, window_size)
    return torch.logsumexp(s, dim=-1)


def dequant_e4m3(cache: torch.Tensor, scale=None) -> torch.Tensor:
    """e4m3 KV cache -> fp32 values: byte value * scale, scale a float, a 1-element tensor or a
    (kv_heads,) tensor applied along the head axis (dim -2).  The oracle of the e4m3 decode is dequantise-then-fp32."""
    x = cache.float()
    if scale is None:
        return x
    s = torch.as_tensor(scale, dtype=torch.float32, device=x.device).reshape(-1)
    return x * (s.view(-1, 1) if s.numel() > 1 else s)


def gather_paged(cache: torch.Tensor, block_table_row: torch.Tensor, length: int) -> torch.Tensor:
    """cache (num_blocks, page, Hkv, D) -> the first `length` logical rows (length, Hkv, D), gathered per row."""
    page = cache.shape[1]
    n_pages = (length + page - 1) // page
    idx = block_table_row[:n_pages].long()
    return cache[idx].reshape(n_pages * page, cache.shape[2], cache.shape[3])[:length]


def varlen_ref(q, k, v, cu_seqlens_q, cu_seqlens_k, causal=False, block_table=None, bottom_right=False,
               window_size=(-1, -1)) -> torch.Tensor:
    """Packed varlen: q (total_q,H,D); k/v (total_k,Hkv,D) or paged (num_blocks,page,Hkv,D)."""
    out = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
    cq = cu_seqlens_q.tolist()
    ck = cu_seqlens_k.tolist()
    for b in range(len(cq) - 1):
        lq, lk = cq[b + 1] - cq[b], ck[b + 1] - ck[b]
        if lq == 0:
            continue
        qb = q[cq[b]:cq[b + 1]][None]
        if block_table is not None:
            kb = gather_paged(k, block_table[b], lk)[None]
            vb = gather_paged(v, block_table[b], lk)[None]
        else:
            kb = k[ck[b]:ck[b + 1]][None]
            vb = v[ck[b]:ck[b + 1]][None]
        if lk == 0:
            continue
        out[cq[b]:cq[b + 1]] = attention_ref(qb, kb, vb, causal=causal, bottom_right=bottom_right,
                                             window_size=window_size)[0]
    return out


def kvcache_ref(q, k_cache, v_cache, cache_seqlens=None, block_table=None, return_lse=False, window_left=-1):
    """q (B,1,H,D); contiguous cache (B,Sk,Hkv,D) or paged (num_blocks,page,Hkv,D) + block_table."""
    B, _, H, D = q.shape
    out = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
    lse = torch.full((B, H), float("-inf"), dtype=torch.float32, device=q.device)
    for b in range(B):
        if block_table is not None:
            cap = block_table.shape[1] * k_cache.shape[1]
        else:
            cap = k_cache.shape[1]
        if cache_seqlens is None:
            L = cap
        elif isinstance(cache_seqlens, int):
            L = cache_seqlens
        else:
            L = int(cache_seqlens[b])
        if L == 0:
            continue
        if block_table is not None:
            kb = gather_paged(k_cache, block_table[b], L)[None]
            vb = gather_paged(v_cache, block_table[b], L)[None]
        else:
            kb = k_cache[b:b + 1, :L]
            vb = v_cache[b:b + 1, :L]
        if window_left >= 0:   # sliding window: the query at position L - 1 sees the last window_left + 1 keys
            kb, vb = kb[:, max(L - 1 - window_left, 0):], vb[:, max(L - 1 - window_left, 0):]
        out[b:b + 1] = attention_ref(q[b:b + 1], kb, vb)
        if return_lse:
            Hkv = kb.shape[2]
            kf = _expand_kv(kb.float(), H // Hkv, 2)[0]            # (L,H,D)
            s = torch.einsum("hd,lhd->hl", q[b, 0].float(), kf) / math.sqrt(D)
            lse[b] = torch.logsumexp(s, dim=-1)
    return (out, lse) if return_lse else out


# --------------------------------------------------------------------------- C oracle
class _OracleParams(ctypes.Structure):
    _fields_ = [
        ("q", ctypes.c_void_p), ("k", ctypes.c_void_p), ("v", ctypes.c_void_p), ("o", ctypes.c_void_p),
        ("q_batch_stride", ctypes.c_int64), ("q_head_stride", ctypes.c_int64), ("q_row_stride", ctypes.c_int64),
        ("k_batch_stride", ctypes.c_int64), ("k_head_stride", ctypes.c_int64), ("k_row_stride", ctypes.c_int64),
        ("v_batch_stride", ctypes.c_int64), ("v_head_stride", ctypes.c_int64), ("v_row_stride", ctypes.c_int64),
        ("o_batch_stride", ctypes.c_int64), ("o_head_stride", ctypes.c_int64), ("o_row_stride", ctypes.c_int64),
        ("batch", ctypes.c_int32), ("heads", ctypes.c_int32), ("kv_heads", ctypes.c_int32),
        ("head_dim", ctypes.c_int32), ("seqlen_q", ctypes.c_int32), ("seqlen_k", ctypes.c_int32),
        ("cu_seqlens_q", ctypes.c_void_p), ("cu_seqlens_k", ctypes.c_void_p),
        ("block_table", ctypes.c_void_p),
        ("block_table_batch_stride", ctypes.c_int32), ("page_block_size", ctypes.c_int32),
        ("k_cache_block_stride", ctypes.c_int64), ("v_cache_block_stride", ctypes.c_int64),
        ("seqlens_k", ctypes.c_void_p),
        ("is_causal", ctypes.c_int32), ("p_round", ctypes.c_int32), ("num_splits", ctypes.c_int32),
        ("lse", ctypes.c_void_p),
    ]


def build_c_oracle(force: bool = False) -> str:
    """gcc -O2 -shared oracle/attention_oracle.c -> oracle/_build/liboracle.so"""
    src = os.path.join(_HERE, "attention_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(_BUILD, exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-o", _SO, src, "-lm"])
    return _SO


_lib = None


def _oracle_lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build_c_oracle())
        _lib.oracle_prefill.argtypes = [ctypes.POINTER(_OracleParams)]
        _lib.oracle_decode.argtypes = [ctypes.POINTER(_OracleParams)]
        assert _lib.oracle_params_size() == ctypes.sizeof(_OracleParams)
    return _lib


def _p_round(dtype) -> int:
    return {torch.float16: 1, torch.bfloat16: 2}.get(dtype, 0)


def _i32(t):
    return None if t is None else t.detach().to("cpu", torch.int32).contiguous()


def c_prefill(q, k, v, causal=False, cu_seqlens_q=None, cu_seqlens_k=None, block_table=None) -> torch.Tensor:
    """C restatement of the reference prefill kernel.  Returns float32 with q's shape."""
    in_dtype = q.dtype
    qf, kf, vf = (t.detach().to("cpu", torch.float32).contiguous() for t in (q, k, v))
    o = torch.zeros_like(qf)
    P = _OracleParams()
    P.q, P.k, P.v, P.o = qf.data_ptr(), kf.data_ptr(), vf.data_ptr(), o.data_ptr()
    keep = [qf, kf, vf, o]
    varlen = cu_seqlens_q is not None
    if varlen:
        cq, ck = _i32(cu_seqlens_q), _i32(cu_seqlens_k)
        keep += [cq, ck]
        P.cu_seqlens_q, P.cu_seqlens_k = cq.data_ptr(), ck.data_ptr()
        P.batch = cq.numel() - 1
        P.heads, P.head_dim = qf.shape[1], qf.shape[2]
        P.q_row_stride, P.q_head_stride = qf.stride(0), qf.stride(1)
        P.o_row_stride, P.o_head_stride = o.stride(0), o.stride(1)
    else:
        P.batch, P.seqlen_q, P.heads, P.head_dim = qf.shape
        P.seqlen_k = kf.shape[1]
        P.q_batch_stride, P.q_row_stride, P.q_head_stride = qf.stride(0), qf.stride(1), qf.stride(2)
        P.o_batch_stride, P.o_row_stride, P.o_head_stride = o.stride(0), o.stride(1), o.stride(2)
        P.k_batch_stride, P.v_batch_stride = kf.stride(0), vf.stride(0)
    P.kv_heads = kf.shape[-2]
    P.k_row_stride, P.k_head_stride = kf.stride(-3), kf.stride(-2)
    P.v_row_stride, P.v_head_stride = vf.stride(-3), vf.stride(-2)
    if block_table is not None:
        bt = _i32(block_table)
        keep.append(bt)
        P.block_table = bt.data_ptr()
        P.block_table_batch_stride = bt.stride(0)
        P.page_block_size = kf.shape[1]
        P.k_cache_block_stride, P.v_cache_block_stride = kf.stride(0), vf.stride(0)
    P.is_causal = int(bool(causal))
    P.p_round = _p_round(in_dtype)
    rc = _oracle_lib().oracle_prefill(ctypes.byref(P))
    assert rc == 0
    return o


def c_decode(q, k_cache, v_cache, cache_seqlens=None, block_table=None, num_splits=1, return_lse=False):
    """C restatement of the reference split-KV decode + combine.  q (B,1,H,D).  Returns float32."""
    qf, kf, vf = (t.detach().to("cpu", torch.float32).contiguous() for t in (q, k_cache, v_cache))
    o = torch.zeros_like(qf)
    B, _, H, D = qf.shape
    lse = torch.empty(B, H, dtype=torch.float32)
    P = _OracleParams()
    P.q, P.k, P.v, P.o = qf.data_ptr(), kf.data_ptr(), vf.data_ptr(), o.data_ptr()
    P.lse = lse.data_ptr()
    keep = [qf, kf, vf, o, lse]
    P.batch, P.seqlen_q, P.heads, P.head_dim = B, 1, H, D
    P.kv_heads = kf.shape[-2]
    P.q_batch_stride, P.q_row_stride, P.q_head_stride = qf.stride(0), qf.stride(1), qf.stride(2)
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = o.stride(0), o.stride(1), o.stride(2)
    P.k_row_stride, P.k_head_stride = kf.stride(-3), kf.stride(-2)
    P.v_row_stride, P.v_head_stride = vf.stride(-3), vf.stride(-2)
    if block_table is not None:
        bt = _i32(block_table)
        keep.append(bt)
        P.block_table = bt.data_ptr()
        P.block_table_batch_stride = bt.stride(0)
        P.page_block_size = kf.shape[1]
        P.k_cache_block_stride, P.v_cache_block_stride = kf.stride(0), vf.stride(0)
        P.seqlen_k = bt.shape[1] * kf.shape[1]
    else:
        P.k_batch_stride, P.v_batch_stride = kf.stride(0), vf.stride(0)
        P.seqlen_k = kf.shape[1]
    if cache_seqlens is not None:
        if isinstance(cache_seqlens, int):
            cache_seqlens = torch.full((B,), cache_seqlens, dtype=torch.int32)
        sl = _i32(cache_seqlens)
        keep.append(sl)
        P.seqlens_k = sl.data_ptr()
    P.num_splits = int(num_splits)
    rc = _oracle_lib().oracle_decode(ctypes.byref(P))
    assert rc == 0
    return (o, lse) if return_lse else o


def merge_partials_ref(o_parts: Sequence[torch.Tensor], lse_parts: Sequence[torch.Tensor]):
    """LSE merge of normalised partials: O = sum_r exp(lse_r - lse) O_r, lse = logsumexp_r lse_r."""
    lse_all = torch.stack([x.float() for x in lse_parts])            # (R, ...)
    lse = torch.logsumexp(lse_all, dim=0)
    w = torch.exp(lse_all - lse.unsqueeze(0))
    w = torch.nan_to_num(w, nan=0.0)
    o = sum(wi.unsqueeze(-1) * oi.float() for wi, oi in zip(w, o_parts))
    return o, lse
