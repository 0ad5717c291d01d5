"""KV-sharded decode with the exchange fused into the merge kernel (csrc/xchg_sm100.cu, include/mfa_b200.h mfa_xchg_*).
On ONE GPU the R ranks live in one process: each has its own gather buffer, the buffers are connected by raw pointers,
and the ranks' kernels run on R different streams -- the same kernels, flags and peer stores as the multi-process
NVLink run (scripts/multi_gpu_check.py covers that one under torchrun).  Oracle: the unsharded fp32 decode."""
import pytest
import torch

from conftest import TOL as _TOL

pytestmark = pytest.mark.gpu
TOL = {torch.float16: _TOL["fp16"], torch.bfloat16: _TOL["bf16"]}


def _problem(B, H, Hkv, D, lens, page, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    cap = max(lens)
    per = (cap + page - 1) // page
    nb = B * per
    kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt, generator=g)
    vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt, generator=g)
    bt = torch.randperm(nb, device="cuda", generator=g).int().reshape(B, per)
    q = torch.randn(B, 1, H, D, device="cuda", dtype=dt, generator=g)
    sl = torch.tensor(lens, dtype=torch.int32, device="cuda")
    return q, kc, vc, bt, sl


def _shard_args(sl, bt, world, rows, page):
    """Per-rank lengths / page tables, computed BEFORE the concurrent region: in ONE process a first-time kernel load
    (CUDA lazy loading) or a cudaMalloc synchronises the context and would wait for the peer kernels that are spinning
    on this rank's push.  (One process per GPU -- the real deployment -- has one context per rank and no such coupling.)"""
    from mini_flash_attention_b200 import parallel as par
    args = [(par.kv_shard_lengths(sl, r, rows), par.kv_shard_block_table(bt, r, rows, page)) for r in range(world)]
    torch.cuda.synchronize()
    return args


def _warm(streams, q, kc, vc, args, splits):
    """Load the kernels and fill each stream's allocator pool with a world-1 exchange on every stream."""
    from mini_flash_attention_b200 import parallel as par
    x1 = par.KVShardExchange(1, 0, q.shape[0] * q.shape[2], q.shape[3])
    for st, (l, t) in zip(streams, args):
        with torch.cuda.stream(st):
            par.decode_kv_sharded(q, kc, vc, l, t, world=1, num_splits=splits, exchange=x1)
        torch.cuda.synchronize()
    x1.close()


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("B,H,Hkv,D,lens,page,dt,splits", [
    (1, 32, 8, 128, [5000], 64, torch.bfloat16, 0),
    (3, 8, 2, 128, [4096, 130, 1], 256, torch.float16, 0),      # ragged: some shards are empty (LSE -inf)
    (2, 16, 16, 64, [777, 2048], 64, torch.float16, 1),         # one split per rank: the decode kernel feeds the merge
    (2, 4, 1, 256, [1500, 900], 64, torch.bfloat16, 3),
])
def test_fused_exchange_matches_unsharded(world, B, H, Hkv, D, lens, page, dt, splits):
    from mini_flash_attention_b200 import parallel as par
    from oracle import ref_attention as R
    q, kc, vc, bt, sl = _problem(B, H, Hkv, D, lens, page, dt, 3)
    ref = R.kvcache_ref(q, kc, vc, sl, bt)
    rows = par.kv_shard_rows(bt.shape[1] * page, world, align=page)
    xs = par.KVShardExchange.local_group(world, B * H, D)
    streams = [torch.cuda.Stream() for _ in range(world)]
    args = _shard_args(sl, bt, world, rows, page)
    _warm(streams, q, kc, vc, args, splits)
    for step in range(3):   # several steps: epochs and the parity double-buffer
        outs = [None] * world
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                outs[r] = par.decode_kv_sharded(q, kc, vc, args[r][0], args[r][1], world=world,
                                                num_splits=splits, exchange=xs[r])
        torch.cuda.synchronize()
        for r in range(world):
            assert xs[r].status() == 0
            err = (outs[r].float() - ref).abs().max().item()
            assert err <= TOL[dt], (step, r, err)
            assert torch.equal(outs[r], outs[0])   # every rank merges the same rows in the same order
    for x in xs:
        x.close()


def test_fused_exchange_equals_allgather_path():
    """Same partials, same merge arithmetic: the fused kernel and pack -> gather -> combine agree to fp32 rounding."""
    from mini_flash_attention_b200 import _C, parallel as par
    world, B, H, Hkv, D, page, dt = 2, 2, 32, 8, 128, 64, torch.bfloat16
    q, kc, vc, bt, sl = _problem(B, H, Hkv, D, [3000, 1234], page, dt, 5)
    rows = par.kv_shard_rows(bt.shape[1] * page, world, align=page)
    packed = []
    for r in range(world):
        o, l = _C.decode_partial(q, kc, vc, par.kv_shard_lengths(sl, r, rows), par.kv_shard_block_table(bt, r, rows, page), 0)
        packed.append(par.pack_partial(o, l))
    base = _C.combine_partials(torch.stack(packed), dt).view(B, 1, H, D)
    xs = par.KVShardExchange.local_group(world, B * H, D)
    streams = [torch.cuda.Stream() for _ in range(world)]
    args = _shard_args(sl, bt, world, rows, page)
    _warm(streams, q, kc, vc, args, 0)
    outs = []
    for r in range(world):
        with torch.cuda.stream(streams[r]):
            outs.append(par.decode_kv_sharded(q, kc, vc, args[r][0], args[r][1], world=world, exchange=xs[r]))
    torch.cuda.synchronize()
    assert all(x.status() == 0 for x in xs)
    assert (outs[0].float() - base.float()).abs().max().item() <= 1e-2 * 0.5
    assert torch.equal(outs[0], outs[1])


def test_missing_peer_times_out_instead_of_hanging():
    """A rank whose peer never launches must come back (bounded spin) and report it through status()."""
    from mini_flash_attention_b200 import parallel as par
    world, B, H, Hkv, D, page, dt = 2, 1, 8, 2, 64, 64, torch.float16
    q, kc, vc, bt, sl = _problem(B, H, Hkv, D, [512], page, dt, 9)
    rows = par.kv_shard_rows(bt.shape[1] * page, world, align=page)
    xs = par.KVShardExchange.local_group(world, B * H, D)
    par.decode_kv_sharded(q, kc, vc, par.kv_shard_lengths(sl, 0, rows), par.kv_shard_block_table(bt, 0, rows, page),
                          world=world, exchange=xs[0])
    torch.cuda.synchronize()
    assert xs[0].status() == 2   # 1 + rank 1
