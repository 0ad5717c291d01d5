"""Single-box multi-GPU partitioner for the attention-forward hot path (one process per GPU).

New work -- the reference (/root/reference) has no distributed code at all; the
arithmetic that makes the KV-axis split legal is its own split-KV + LSE combine
(/root/reference/csrc/mfa/decode.cuh:21-30 partition, :665-755 merge), lifted from "CTAs of one GPU" to
"GPUs of one box".

Two partitions:

* prefill -- independent units per (batch, kv-head group): ``prefill_partition`` cuts the batch first and then
  KV-head groups, every rank runs the single-GPU kernel on its views, NO collective.  The output stays sharded
  along the same dims (``gather_prefill_output`` exists for callers that want it whole; it is not on the timed
  path).
* long-KV decode -- the KV axis is cut into contiguous row ranges, one per rank.  Every rank computes a
  normalised fp32 partial O and its LSE over its own rows (itself split-KV inside the GPU), packs them as one
  (B*Hq, D+1) fp32 buffer, ONE ``all_gather_into_tensor`` over NCCL/NVLink moves all of them (16.5 KB per rank at
  B=1, Hq=32, D=128 -- latency-bound, not bandwidth-bound), and the library's combine kernel merges the R
  partials on every rank.

The compute calls go through ``_C`` (the C-ABI shim); ``partial_fn`` / ``combine_fn`` are injectable so that the
host logic (partition arithmetic, packing, the collective, the merge order) is testable with the gloo backend
on a CPU box where no kernel can run.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import torch


# ------------------------------------------------------------------------------------------- prefill
@dataclass(frozen=True)
class PrefillShard:
    """The (batch, kv-head) box owned by one rank."""
    batch_lo: int
    batch_hi: int
    kv_head_lo: int
    kv_head_hi: int

    def q_heads(self, group: int) -> Tuple[int, int]:
        return self.kv_head_lo * group, self.kv_head_hi * group

    @property
    def empty(self) -> bool:
        return self.batch_hi <= self.batch_lo or self.kv_head_hi <= self.kv_head_lo


def _split_even(n: int, parts: int, idx: int) -> Tuple[int, int]:
    """[lo, hi) of chunk idx when n items are cut into `parts` contiguous chunks whose sizes differ by <= 1."""
    base, rem = divmod(n, parts)
    lo = idx * base + min(idx, rem)
    return lo, lo + base + (1 if idx < rem else 0)


def prefill_partition(batch: int, kv_heads: int, world: int) -> List[PrefillShard]:
    """Cut (batch x kv_heads) units over `world` ranks: batch first; when world > batch the ranks that share a
    batch slice cut the KV heads (q heads follow their KV head, so K/V shards stay disjoint -- SURVEY.md 8e)."""
    if world <= 0:
        raise ValueError("world must be positive")
    if world <= batch:
        return [PrefillShard(*_split_even(batch, world, r), 0, kv_heads) for r in range(world)]
    # more ranks than batch elements: the ranks are cut into `batch` contiguous runs; the ranks of one run
    # share a batch element and cut its KV heads
    shards = []
    for b in range(batch):
        lo, hi = _split_even(world, batch, b)
        for r in range(lo, hi):
            shards.append(PrefillShard(b, b + 1, *_split_even(kv_heads, hi - lo, r - lo)))
    return shards


def prefill_sharded(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = False, *, rank: int, world: int,
                    forward_fn: Optional[Callable] = None) -> Tuple[Optional[torch.Tensor], PrefillShard]:
    """Run this rank's share of flash_attn_func(q, k, v, causal).  q (B,Sq,H,D), k/v (B,Sk,Hkv,D) are the FULL
    problem (or any tensor with the full shape -- only this rank's views are read).  Returns (out_shard, shard)
    with out_shard of shape (b_hi-b_lo, Sq, (h_hi-h_lo)*group, D), or (None, shard) for a rank with no work.
    No collective."""
    if forward_fn is None:
        from . import _C
        forward_fn = lambda a, b, c, cz: _C.mini_flash_attention_forward(a, b, c, None, cz, -1, -1)  # noqa: E731
    B, _, H, _ = q.shape
    Hkv = k.shape[2]
    if H % Hkv:
        raise RuntimeError("number of key/value heads must be divisible by number of query heads")
    group = H // Hkv
    shard = prefill_partition(B, Hkv, world)[rank]
    if shard.empty:
        return None, shard
    h_lo, h_hi = shard.q_heads(group)
    qs = q[shard.batch_lo:shard.batch_hi, :, h_lo:h_hi]
    ks = k[shard.batch_lo:shard.batch_hi, :, shard.kv_head_lo:shard.kv_head_hi]
    vs = v[shard.batch_lo:shard.batch_hi, :, shard.kv_head_lo:shard.kv_head_hi]
    return forward_fn(qs, ks, vs, causal), shard


def gather_prefill_output(out_shard: Optional[torch.Tensor], shard: PrefillShard, full_shape, dtype, device, *, world: int,
                          group=None) -> torch.Tensor:
    """Assemble the full (B,Sq,H,D) output on every rank (convenience; NOT on the timed path)."""
    import torch.distributed as dist
    B, Sq, H, D = full_shape
    full = torch.zeros(full_shape, dtype=dtype, device=device)
    if out_shard is not None:
        h_per_kv = out_shard.shape[2] // (shard.kv_head_hi - shard.kv_head_lo)
        full[shard.batch_lo:shard.batch_hi, :, shard.kv_head_lo * h_per_kv:shard.kv_head_hi * h_per_kv] = out_shard
    if world > 1:
        # shards are disjoint boxes, so a sum over ranks of zero-padded tensors is the concatenation
        work = full.float()
        dist.all_reduce(work, group=group)
        full = work.to(dtype)
    return full


# ------------------------------------------------------------------------------------------- decode
def kv_shard_rows(capacity_rows: int, world: int, align: int = 64) -> int:
    """Rows of the KV axis owned by each rank: ceil(capacity / world) rounded up to `align` (page size or the
    kernel's 64-row tile, so that shard boundaries never cut a page)."""
    per = -(-capacity_rows // world)
    return -(-per // align) * align


def kv_shard_lengths(cache_seqlens: torch.Tensor, rank: int, rows_per_rank: int) -> torch.Tensor:
    """Local lengths of rank `rank`: clamp(len - rank*rows_per_rank, 0, rows_per_rank)  (SURVEY.md 8e)."""
    return (cache_seqlens.to(torch.int64) - rank * rows_per_rank).clamp_(0, rows_per_rank).to(torch.int32)


def kv_shard_block_table(block_table: torch.Tensor, rank: int, rows_per_rank: int, page: int) -> torch.Tensor:
    """This rank's slice of the page table (columns [rank*rows/page, (rank+1)*rows/page))."""
    if rows_per_rank % page:
        raise ValueError("rows_per_rank must be a multiple of the page size")
    per = rows_per_rank // page
    sl = block_table[:, rank * per:(rank + 1) * per]
    if sl.shape[1] < per:  # ragged tail: pad with page 0 (never read, lengths are clamped)
        pad = torch.zeros(sl.shape[0], per - sl.shape[1], dtype=sl.dtype, device=sl.device)
        sl = torch.cat([sl, pad], dim=1)
    return sl.contiguous()


def pack_partial(o_part: torch.Tensor, lse_part: torch.Tensor) -> torch.Tensor:
    """(B,H,D) fp32 + (B,H) fp32 -> (B*H, D+1) fp32, per row [O_0..O_{D-1}, LSE]: one buffer, one collective."""
    B, H, D = o_part.shape
    packed = torch.empty(B * H, D + 1, dtype=torch.float32, device=o_part.device)
    packed[:, :D] = o_part.reshape(B * H, D)
    packed[:, D] = lse_part.reshape(B * H)
    return packed


class KVShardExchange:
    """Peer-memory exchange buffers of the KV-sharded decode (include/mfa_b200.h `mfa_xchg_*`): one gather buffer per
    rank, mapped into every peer with CUDA IPC, so that the merge kernel of `decode_kv_sharded(..., exchange=x)` pushes
    its partial row straight into the peers' memory over NVLink instead of going through an NCCL all-gather.

    One process per GPU: `KVShardExchange(world, rank, max_rows, head_dim, group=g)` allocates the local buffer on the
    current device, all-gathers the 64-byte IPC handles over `group` (torch.distributed, any backend) and opens the
    peers.  Collective: every rank of the group must construct it, and later call decode_kv_sharded the same number of
    times in the same order.  `KVShardExchange.local_group(world, ...)` builds all ranks inside ONE process on one
    device (tests: the ranks' kernels then run on different streams of the same GPU)."""

    def __init__(self, world: int, rank: int, max_rows: int, head_dim: int, group=None, _connect: bool = True):
        import ctypes
        from . import _lib
        self._lib = _lib.load()
        self.world, self.rank, self.max_rows, self.head_dim = world, rank, max_rows, head_dim
        self.verify_steps, self._steps = 2, 0   # decode_kv_sharded checks the timeout status on the first calls
        h = ctypes.c_void_p()
        _lib.check(self._lib.mfa_xchg_create(world, rank, max_rows, head_dim, ctypes.byref(h)), "mfa_xchg_create")
        self.handle = h.value
        if _connect and world > 1:
            import torch.distributed as dist
            mine = ctypes.create_string_buffer(_lib.XCHG_HANDLE_BYTES)
            _lib.check(self._lib.mfa_xchg_export(self.handle, mine), "mfa_xchg_export")
            handles = [None] * world
            dist.all_gather_object(handles, bytes(mine.raw), group=group)
            blob = ctypes.create_string_buffer(b"".join(handles), world * _lib.XCHG_HANDLE_BYTES)
            _lib.check(self._lib.mfa_xchg_connect(self.handle, blob), "mfa_xchg_connect")
            dist.barrier(group=group)   # every rank has mapped every buffer before the first push

    @classmethod
    def local_group(cls, world: int, max_rows: int, head_dim: int):
        import ctypes
        from . import _lib
        xs = [cls(world, r, max_rows, head_dim, _connect=False) for r in range(world)]
        ptrs = (ctypes.c_void_p * world)(*[x._lib.mfa_xchg_local_ptr(x.handle) for x in xs])
        for x in xs:
            x.verify_steps = 0   # the ranks are streams of ONE host thread: a status read (device sync) between their launches would deadlock
            if world > 1:
                _lib.check(x._lib.mfa_xchg_connect_ptrs(x.handle, ptrs), "mfa_xchg_connect_ptrs")
        return xs

    def status(self, clear: bool = False) -> int:
        """0, or 1 + the rank whose push did not arrive within the kernel's timeout (~2 s; MFA_XCHG_TIMEOUT_MS).
        Synchronises the device.  Sticky until read with clear=True.  Rows that timed out are NaN in the output."""
        fn = self._lib.mfa_xchg_clear_status if clear else self._lib.mfa_xchg_status
        return int(fn(self.handle))

    def check(self) -> None:
        """Raise if any step since the last check timed out waiting for a peer (synchronises the device)."""
        st = self.status(clear=True)
        if st:
            raise RuntimeError(f"KV-shard exchange: the push of rank {st - 1} did not arrive within the timeout "
                               f"(rank {self.rank} of {self.world}); the affected output rows are NaN")

    def close(self) -> None:
        if getattr(self, "handle", None):
            self._lib.mfa_xchg_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def decode_kv_sharded(q: torch.Tensor, k_cache_local: torch.Tensor, v_cache_local: torch.Tensor,
                      local_seqlens: torch.Tensor, block_table_local: Optional[torch.Tensor] = None, *,
                      world: int, group=None, num_splits: int = 0, partial_fn: Optional[Callable] = None,
                      combine_fn: Optional[Callable] = None, gathered: Optional[torch.Tensor] = None,
                      exchange: Optional[KVShardExchange] = None) -> torch.Tensor:
    """flash_attn_with_kvcache with the KV axis sharded over `world` ranks.

    q (B,1,Hq,D) is replicated; k_cache_local / v_cache_local / block_table_local / local_seqlens describe only
    THIS rank's rows of every sequence (see kv_shard_lengths / kv_shard_block_table).  Returns the full result
    (B,1,Hq,D) in q's dtype on every rank.  One all-gather of B*Hq*(D+1) floats per rank -- or, with
    `exchange` (a KVShardExchange), no collective at all: the merge kernel pushes the rows over NVLink peer memory."""
    if exchange is not None:
        from . import _C
        if exchange.world != world:
            raise ValueError("exchange was built for another world size")
        out = _C.decode_kv_sharded_fused(q, k_cache_local, v_cache_local, local_seqlens, block_table_local, num_splits,
                                         exchange.handle)
        # Contract: a peer that does not arrive within the timeout leaves NaN rows and a sticky status word.  The first
        # `verify_steps` calls (module loads, late ranks: the risky ones) check it here, at the cost of a device
        # synchronisation; afterwards the caller polls exchange.check() at its own pace (never inside a graph capture).
        if exchange._steps < exchange.verify_steps and not torch.cuda.is_current_stream_capturing():
            exchange.check()
        exchange._steps += 1
        return out
    if partial_fn is None or combine_fn is None:
        from . import _C
        partial_fn = partial_fn or _C.decode_partial
        combine_fn = combine_fn or _C.combine_partials
    B, one, H, D = q.shape
    o_part, lse_part = partial_fn(q, k_cache_local, v_cache_local, local_seqlens, block_table_local, num_splits)
    packed = pack_partial(o_part, lse_part)
    if world > 1:
        import torch.distributed as dist
        if gathered is None:
            gathered = torch.empty(world, B * H, D + 1, dtype=torch.float32, device=q.device)
        dist.all_gather_into_tensor(gathered.view(world * B * H, D + 1), packed, group=group)
    else:
        gathered = packed.unsqueeze(0)
    out = combine_fn(gathered, q.dtype)
    return out.view(B, 1, H, D)
