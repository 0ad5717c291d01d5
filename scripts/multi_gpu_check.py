#!/usr/bin/env python
"""Multi-GPU leg of the hot path on real GPUs, one process per GPU over NCCL:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        scripts/multi_gpu_check.py [--quick]

1. parity: KV-sharded decode against the fp32 oracle on the full, unsharded problem (paged cache, ragged lengths,
   GQA), both forms: partial O + LSE per rank -> ONE all_gather -> combine kernel, and the exchange fused into the
   merge kernel over NVLink peer memory (KVShardExchange: no NCCL call on the data path);
2. parity: prefill sharded by (batch, kv-head) with no collective, every rank's shard against the oracle;
3. timing (device events, max over ranks): long-context decode KV = 128K ... 1M rows split over the ranks,
   B=1, Hq=32, Hkv=8, D=128, and head-sharded prefill S=16K.
Rank 0 prints one JSON line per measurement (also appended to gpurun_out/multigpu.jsonl when that directory exists).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def emit(rank, rec):
    if rank != 0:
        return
    line = json.dumps(rec)
    print(line, flush=True)
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "multigpu.jsonl"), "a") as f:
            f.write(line + "\n")


def timed(fn, steps, warmup, world):
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def latency(fn, steps, warmup, world):
    """One ISOLATED step at a time (ranks aligned by a barrier + device sync before each): launch-to-completion time of
    a single decode step, median over `steps`, max over ranks -- what a serving loop that needs the token waits for."""
    ts = []
    for i in range(warmup + steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            ts.append(e0.elapsed_time(e1))
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--parity-only", action="store_true", help="sections 1 and 2 only (tests/test_multigpu.py)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import mini_flash_attention as mfa  # noqa: F401
    from mini_flash_attention_b200 import parallel as P
    from oracle import ref_attention as R

    dt = torch.bfloat16
    tol = 1e-2
    # ------------------------------------------------------------------ 1. KV-sharded decode parity
    g = torch.Generator(device="cuda").manual_seed(1234)      # same seed on every rank: same full problem
    B, Hq, Hkv, D, page = 3, 16, 4, 128, 64
    cap = 2048                                                 # rows of capacity per sequence
    nblk = B * cap // page
    q = torch.randn(B, 1, Hq, D, device="cuda", dtype=dt, generator=g)
    kc = torch.randn(nblk, page, Hkv, D, device="cuda", dtype=dt, generator=g)
    vc = torch.randn(nblk, page, Hkv, D, device="cuda", dtype=dt, generator=g)
    bt = torch.randperm(nblk, device="cuda", generator=g).reshape(B, cap // page).int()
    lens = torch.tensor([2048, 700, 1], dtype=torch.int32, device="cuda")
    rows = P.kv_shard_rows(cap, world, align=page)
    out = P.decode_kv_sharded(q, kc, vc, P.kv_shard_lengths(lens, rank, rows), P.kv_shard_block_table(bt, rank, rows, page),
                              world=world)
    ref = R.kvcache_ref(q, kc, vc, lens, bt)
    err = (out.float() - ref).abs().max()
    if world > 1:
        dist.all_reduce(err, op=dist.ReduceOp.MAX)
    emit(rank, {"check": "decode_kv_sharded parity", "world": world, "max_abs_err": float(err), "tol": tol,
                "ok": bool(float(err) <= tol)})
    assert float(err) <= tol, float(err)
    # the same through the fused exchange (CUDA IPC peer buffers, push + flag + merge in one kernel)
    xchg = P.KVShardExchange(world, rank, B * Hq, D)
    sl_r, bt_r = P.kv_shard_lengths(lens, rank, rows), P.kv_shard_block_table(bt, rank, rows, page)
    for step in range(3):
        out_f = P.decode_kv_sharded(q, kc, vc, sl_r, bt_r, world=world, exchange=xchg)
    st = torch.tensor([float(xchg.status())], device="cuda")
    ferr = (out_f.float() - ref).abs().max()
    same = (out_f.float() - out.float()).abs().max()
    if world > 1:
        dist.all_reduce(ferr, op=dist.ReduceOp.MAX)
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        dist.all_reduce(same, op=dist.ReduceOp.MAX)
    emit(rank, {"check": "decode_kv_sharded FUSED exchange parity", "world": world, "max_abs_err": float(ferr), "tol": tol,
                "vs_allgather_path": float(same), "timeout_status": int(st.item()),
                "ok": bool(float(ferr) <= tol and int(st.item()) == 0)})
    assert float(ferr) <= tol and int(st.item()) == 0, (float(ferr), int(st.item()))
    xchg.close()

    # ------------------------------------------------------------------ 2. sharded prefill parity
    Bp, Sp, Hp, Hkp = 2, 384, 8, 2
    qp = torch.randn(Bp, Sp, Hp, D, device="cuda", dtype=dt, generator=g)
    kp = torch.randn(Bp, Sp, Hkp, D, device="cuda", dtype=dt, generator=g)
    vp = torch.randn(Bp, Sp, Hkp, D, device="cuda", dtype=dt, generator=g)
    o_sh, shard = P.prefill_sharded(qp, kp, vp, True, rank=rank, world=world)
    full_ref = R.attention_ref(qp, kp, vp, True)
    perr = torch.zeros((), device="cuda")
    if o_sh is not None:
        grp = Hp // Hkp
        h_lo, h_hi = shard.q_heads(grp)
        perr = (o_sh.float() - full_ref[shard.batch_lo:shard.batch_hi, :, h_lo:h_hi]).abs().max()
    if world > 1:
        dist.all_reduce(perr, op=dist.ReduceOp.MAX)
    emit(rank, {"check": "prefill_sharded parity", "world": world, "max_abs_err": float(perr), "tol": tol,
                "ok": bool(float(perr) <= tol)})
    assert float(perr) <= tol, float(perr)

    if args.parity_only:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    # ------------------------------------------------------------------ 3. timings
    Bl, Hq, Hkv, page = 1, 32, 8, 256
    kv_list = [131072] if args.quick else [131072, 262144, 524288, 1048576]
    for KV in kv_list:
        rows = P.kv_shard_rows(KV, world, align=page)
        nb = Bl * rows // page
        ql = torch.randn(Bl, 1, Hq, D, device="cuda", dtype=dt, generator=g)
        kl = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        vl = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        btl = torch.randperm(nb, device="cuda").reshape(Bl, rows // page).int()
        full_len = torch.full((Bl,), KV, dtype=torch.int32, device="cuda")
        ll = P.kv_shard_lengths(full_len, rank, rows)
        gathered = torch.empty(world, Bl * Hq, D + 1, dtype=torch.float32, device="cuda")
        fn = lambda: P.decode_kv_sharded(ql, kl, vl, ll, btl, world=world, gathered=gathered)  # noqa: E731
        ms = timed(fn, 30, 5, world)
        nbytes = 2.0 * Bl * KV * Hkv * D * 2
        emit(rank, {"bench": "long-context decode, KV axis sharded", "world": world, "kv": KV, "B": Bl, "Hq": Hq, "Hkv": Hkv,
                    "ms": round(ms, 4), "aggregate_GBps": round(nbytes / ms / 1e6, 1),
                    "per_gpu_GBps": round(nbytes / ms / 1e6 / world, 1), "collective": "1 all_gather of B*Hq*(D+1) fp32 per rank"})
        xl = P.KVShardExchange(world, rank, Bl * Hq, D)
        fn2 = lambda: P.decode_kv_sharded(ql, kl, vl, ll, btl, world=world, exchange=xl)  # noqa: E731
        ms2 = timed(fn2, 30, 5, world)
        lat_ag = latency(fn, 20, 3, world)
        lat_fused = latency(fn2, 20, 3, world)
        # the fused step has launch parameters that never change (the epoch lives on the device): capture + replay
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                fn2()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out_g = fn2()
        ms3 = timed(graph.replay, 30, 5, world)
        lat_graph = latency(graph.replay, 20, 3, world)
        ref_out = fn()
        torch.cuda.synchronize()
        gerr = (out_g.float() - ref_out.float()).abs().max()
        if world > 1:
            dist.all_reduce(gerr, op=dist.ReduceOp.MAX)
        st = xl.status()
        emit(rank, {"bench": "long-context decode, KV axis sharded, FUSED exchange", "world": world, "kv": KV, "B": Bl, "Hq": Hq,
                    "Hkv": Hkv, "ms": round(ms2, 4), "ms_graph_replay": round(ms3, 4),
                    "aggregate_GBps": round(nbytes / ms2 / 1e6, 1), "aggregate_GBps_graph": round(nbytes / ms3 / 1e6, 1),
                    "per_gpu_GBps_graph": round(nbytes / ms3 / 1e6 / world, 1),
                    "step_latency_ms": {"allgather_path": round(lat_ag, 4), "fused": round(lat_fused, 4),
                                        "fused_graph": round(lat_graph, 4)},
                    "graph_vs_allgather_max_abs_diff": float(gerr), "timeout_status": st,
                    "collective": "none: P2P push + flag inside the merge kernel (NVLink peer memory)"})
        del graph, out_g
        xl.close()
        del kl, vl
        torch.cuda.empty_cache()
    # head-sharded prefill S=16K (B=1, H=32, Hkv=8): strong scaling, no collective
    S = 4096 if args.quick else 16384
    qh = torch.randn(1, S, 32, D, device="cuda", dtype=dt)
    kh = torch.randn(1, S, 8, D, device="cuda", dtype=dt)
    vh = torch.randn(1, S, 8, D, device="cuda", dtype=dt)
    for causal in (False, True):
        ms = timed(lambda: P.prefill_sharded(qh, kh, vh, causal, rank=rank, world=world), 5, 2, world)
        fl = 4.0 * 32 * S * S * D * (0.5 if causal else 1.0)
        emit(rank, {"bench": "prefill sharded by kv-head, no collective", "world": world, "S": S, "H": 32, "Hkv": 8, "causal": causal,
                    "ms": round(ms, 3), "aggregate_TFLOPs": round(fl / ms / 1e9, 1)})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
