"""Host-buffer entry point: flash_attn_func on tensors that live in (pinned) HOST memory.

The reference has no such call -- its extension takes CUDA tensors only (/root/reference/csrc/mfa/api.cpp:131-133
CHECK_DEVICE), so a caller with host data does `q.cuda(); k.cuda(); v.cuda(); out = flash_attn_func(...); out.cpu()`:
three copies in, the kernel, one copy out, back to back.  At the headline shape (B=48, S=4096, H=24, D=128, bf16:
3.6 GB in, 1.2 GB out) that sequence is bound by PCIe (~94 ms against 9 ms of kernel).  This module is the same call
made as a PIPELINE over batch chunks on three streams, so that the link runs in both directions at once and the kernel
hides under the copies:

    copy-in stream :  H2D q,k,v chunk i+1   |  compute stream :  prefill kernel chunk i   |  copy-out stream :  D2H out chunk i-1

The arithmetic is the unmodified public op (`_C.mini_flash_attention_forward` with `out=` into a staging buffer) on
each chunk; chunks are whole batch elements, which are independent, so the result is bit-identical to the
one-shot call.  PyTorch supplies streams, events and pinned/device memory -- plumbing only.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _C


class _Slots:
    """Device staging for `depth` chunks in flight: q, k, v, o per slot, plus the streams and events of the pipeline."""

    def __init__(self, device, dtype, q_shape, k_shape, depth: int):
        self.depth = depth
        self.q = [torch.empty(q_shape, dtype=dtype, device=device) for _ in range(depth)]
        self.k = [torch.empty(k_shape, dtype=dtype, device=device) for _ in range(depth)]
        self.v = [torch.empty(k_shape, dtype=dtype, device=device) for _ in range(depth)]
        self.o = [torch.empty(q_shape, dtype=dtype, device=device) for _ in range(depth)]
        self.s_in = torch.cuda.Stream(device)
        self.s_compute = torch.cuda.Stream(device)
        self.s_out = torch.cuda.Stream(device)


_slots: Dict[Tuple, _Slots] = {}


def flash_attn_func_host(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = False, *,
                         out: Optional[torch.Tensor] = None, device=None, chunk_batch: Optional[int] = None,
                         depth: int = 3) -> torch.Tensor:
    """flash_attn_func for HOST tensors.  q (B,Sq,H,D), k/v (B,Sk,Hkv,D) fp16/bf16 in host memory (pinned memory makes the
    copies asynchronous; pageable memory works but serialises them); returns `out` (B,Sq,H,D) in pinned host memory.

    The call is asynchronous with respect to the host like any CUDA op: the current stream of `device` waits for the last
    copy-out, so `torch.cuda.current_stream().synchronize()` (or an event recorded after the call) means `out` is ready.
    chunk_batch: batch elements per pipeline stage (default: about 1/16 of the batch; scripts/host_chunks.py); depth: chunks in flight."""
    if q.is_cuda or k.is_cuda or v.is_cuda:
        raise RuntimeError("flash_attn_func_host takes host tensors; use flash_attn_func for CUDA tensors")
    if q.dim() != 4 or k.dim() != 4 or v.dim() != 4:
        raise RuntimeError("q, k, v must be 4-D (batch, seqlen, heads, head_dim)")
    if not (q.dtype in (torch.float16, torch.bfloat16) and k.dtype == q.dtype and v.dtype == q.dtype):
        raise RuntimeError("FlashAttention only support fp16 and bf16 data type")
    B = q.shape[0]
    if k.shape[0] != B or v.shape != k.shape:
        raise RuntimeError("batch size of q, k and v must be the same and k, v must have the same shape")
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if out is None:
        out = torch.empty(q.shape, dtype=q.dtype).pin_memory()
    elif out.is_cuda or out.shape != q.shape or out.dtype != q.dtype:
        raise RuntimeError("out must be a host tensor with q's shape and dtype")
    if B == 0 or q.numel() == 0:
        return out
    cb = chunk_batch if chunk_batch else max(1, (B + 15) // 16)
    cb = min(cb, B)
    key = (device, q.dtype, (cb,) + tuple(q.shape[1:]), (cb,) + tuple(k.shape[1:]), depth)
    sl = _slots.get(key)
    if sl is None:
        _slots.clear()   # one staging set at a time: a new shape replaces the old buffers
        sl = _slots[key] = _Slots(device, q.dtype, key[2], key[3], depth)
    cur = torch.cuda.current_stream(device)
    for s in (sl.s_in, sl.s_compute, sl.s_out):
        s.wait_stream(cur)
    n_chunks = (B + cb - 1) // cb
    ev_in = [None] * n_chunks
    ev_done = [None] * n_chunks
    ev_out = [None] * n_chunks
    for i in range(n_chunks):
        b0, b1 = i * cb, min(B, (i + 1) * cb)
        n, slot = b1 - b0, i % depth
        with torch.cuda.stream(sl.s_in):
            if i >= depth:
                sl.s_in.wait_event(ev_done[i - depth])    # the kernel that read this slot's inputs has finished
            sl.q[slot][:n].copy_(q[b0:b1], non_blocking=True)
            sl.k[slot][:n].copy_(k[b0:b1], non_blocking=True)
            sl.v[slot][:n].copy_(v[b0:b1], non_blocking=True)
            ev_in[i] = sl.s_in.record_event()
        with torch.cuda.stream(sl.s_compute):
            sl.s_compute.wait_event(ev_in[i])
            if i >= depth:
                sl.s_compute.wait_event(ev_out[i - depth])  # this slot's previous output has left the device
            _C.mini_flash_attention_forward(sl.q[slot][:n], sl.k[slot][:n], sl.v[slot][:n], sl.o[slot][:n], causal, -1, -1)
            ev_done[i] = sl.s_compute.record_event()
        with torch.cuda.stream(sl.s_out):
            sl.s_out.wait_event(ev_done[i])
            out[b0:b1].copy_(sl.o[slot][:n], non_blocking=True)
            ev_out[i] = sl.s_out.record_event()
    cur.wait_stream(sl.s_out)
    cur.wait_stream(sl.s_in)
    cur.wait_stream(sl.s_compute)
    return out


def release_staging() -> None:
    """Free the cached device staging buffers."""
    _slots.clear()
