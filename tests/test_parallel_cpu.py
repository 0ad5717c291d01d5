"""Host-side logic of the multi-GPU partitioner, on CPU: partition arithmetic, and the N>1 data path
(pack -> one all-gather -> merge) with world_size 2 over the gloo backend.  No kernel can run here, so the two
compute calls are injected from the oracle (tests may do that; the product defaults to the C-ABI)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mini_flash_attention_b200 import parallel as P
from oracle import ref_attention as R


@pytest.mark.parametrize("B,Hkv,W", [(48, 24, 1), (48, 24, 8), (2, 8, 8), (1, 8, 8), (3, 8, 8), (1, 8, 2), (5, 4, 4), (1, 2, 4)])
def test_prefill_partition_is_an_exact_cover(B, Hkv, W):
    shards = P.prefill_partition(B, Hkv, W)
    assert len(shards) == W
    seen = set()
    for s in shards:
        for b in range(s.batch_lo, s.batch_hi):
            for h in range(s.kv_head_lo, s.kv_head_hi):
                assert (b, h) not in seen
                seen.add((b, h))
    assert len(seen) == B * Hkv
    sizes = [(s.batch_hi - s.batch_lo) * (s.kv_head_hi - s.kv_head_lo) for s in shards]
    if W <= B:
        assert max(sizes) - min(sizes) <= Hkv   # batch chunks differ by at most one element


def test_kv_shard_arithmetic():
    lens = torch.tensor([0, 1, 100, 256, 257, 1000], dtype=torch.int32)
    rows = P.kv_shard_rows(1000, 4, align=64)
    assert rows == 256
    total = sum(P.kv_shard_lengths(lens, r, rows) for r in range(4))
    assert torch.equal(total, lens)
    assert torch.equal(P.kv_shard_lengths(lens, 1, rows), torch.tensor([0, 0, 0, 0, 1, 256], dtype=torch.int32))
    bt = torch.arange(2 * 7, dtype=torch.int32).reshape(2, 7)
    s0 = P.kv_shard_block_table(bt, 0, 256, 64)
    s1 = P.kv_shard_block_table(bt, 1, 256, 64)
    assert torch.equal(s0, bt[:, :4]) and s1.shape == (2, 4) and torch.equal(s1[:, :3], bt[:, 4:7])


def test_prefill_sharded_views_match_full_problem():
    torch.manual_seed(0)
    q = torch.randn(2, 33, 8, 32).half()
    k = torch.randn(2, 40, 4, 32).half()
    v = torch.randn(2, 40, 4, 32).half()
    full = R.attention_ref(q, k, v, True)
    fwd = lambda a, b, c, cz: R.attention_ref(a, b, c, cz)  # noqa: E731
    W = 4
    out = torch.zeros_like(full)
    for r in range(W):
        o, s = P.prefill_sharded(q, k, v, True, rank=r, world=W, forward_fn=fwd)
        h_lo, h_hi = s.q_heads(2)
        out[s.batch_lo:s.batch_hi, :, h_lo:h_hi] = o
    assert torch.allclose(out, full, atol=1e-6)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_partial(q, kc, vc, lens, bt, num_splits):
    o, lse = R.kvcache_ref(q, kc, vc, lens, bt, return_lse=True)
    return o[:, 0].contiguous(), lse


def _oracle_combine(gathered, dtype):
    D = gathered.shape[-1] - 1
    o, _ = R.merge_partials_ref([g[:, :D] for g in gathered], [g[:, D] for g in gathered])
    return o.to(dtype)


def _worker(rank, world, port, paged, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(1234)          # same inputs on every rank
        B, H, Hkv, D, page = 3, 8, 2, 64, 64
        cap = 640
        lens = torch.tensor([640, 70, 300], dtype=torch.int32)
        q = torch.randn(B, 1, H, D, generator=g).half()
        if paged:
            nb = B * cap // page
            kc = torch.randn(nb, page, Hkv, D, generator=g).half()
            vc = torch.randn(nb, page, Hkv, D, generator=g).half()
            bt = torch.randperm(nb, generator=g).reshape(B, cap // page).int()
            full = R.kvcache_ref(q, kc, vc, lens, bt)
            rows = P.kv_shard_rows(cap, world, align=page)
            out = P.decode_kv_sharded(q, kc, vc, P.kv_shard_lengths(lens, rank, rows),
                                      P.kv_shard_block_table(bt, rank, rows, page), world=world,
                                      partial_fn=_oracle_partial, combine_fn=_oracle_combine)
        else:
            kc = torch.randn(B, cap, Hkv, D, generator=g).half()
            vc = torch.randn(B, cap, Hkv, D, generator=g).half()
            full = R.kvcache_ref(q, kc, vc, lens, None)
            rows = P.kv_shard_rows(cap, world, align=64)
            lo = rank * rows
            out = P.decode_kv_sharded(q, kc[:, lo:lo + rows].contiguous(), vc[:, lo:lo + rows].contiguous(),
                                      P.kv_shard_lengths(lens, rank, rows), None, world=world,
                                      partial_fn=_oracle_partial, combine_fn=_oracle_combine)
        err = (out.float() - full).abs().max().item()
        ret[rank] = err
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("paged", [False, True])
def test_decode_kv_sharded_world2_gloo(paged):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), paged, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        assert ret[r] < 1e-3, f"rank {r}: {ret[r]}"
