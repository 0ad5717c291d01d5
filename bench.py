#!/usr/bin/env python
"""bench.py -- the hot-path benchmark (driver contract).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json `metric`, configs[1]): prefill flash_attn_func, head_dim 128, B=48, H=24, S=4096,
non-causal, bf16, one "step" = one forward over that batch.  value = whole-job TFLOP/s with inputs resident in
HBM (4*B*H*S^2*D flops); e2e = the same metric through the public API with HOST (pinned) buffers,
H2D of q,k,v and D2H of the output inside the timed region.  The decode leg of the metric (B=24, H=24, D=128,
KV=8192, paged, GB/s vs HBM) is measured in the same run and reported under "decode".

N > 1 (torchrun, one rank per GPU): the batch shards across ranks with no collective (prefill partitions by
(batch, head)) -- every rank runs the per-GPU workload, scaling "weak", value = sum of flops /
max-over-ranks time.

--impl reference times the UNMODIFIED reference CUDA extension (oracle/_ref, built from /root/reference by
oracle/build_ref.py) on the same config; when that .so cannot be loaded it falls back to the CPU oracle port.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "prefill TFLOP/s (hd128,B48,H24,S4096) vs bf16 peak; decode KV GB/s vs HBM"
WORKLOAD = dict(B=48, H=24, S=4096, D=128)
DECODE = dict(B=24, H=24, D=128, KV=8192, page=256)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(tflops=d["bf16_tflops"], tflops_sustained=d.get("bf16_tflops_sustained"), hbm=d["hbm_gbs"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(tflops=1590.0, tflops_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """Samples SM clock and throttle reasons DURING the timed region (pynvml, else nvidia-smi)."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.max_mhz, self.power = index, [], set(), None, []
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {}
        if nv is not None:
            for n in ("HwSlowdown", "HwThermalSlowdown", "SwThermalSlowdown", "SwPowerCap", "HwPowerBrakeSlowdown"):
                v = getattr(nv, "nvmlClocksEventReason" + n, None) or getattr(nv, "nvmlClocksThrottleReason" + n, None)
                if v is not None:
                    names[v] = n
        while not self._stop.is_set():
            try:
                if nv is not None:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    try:
                        self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                    except Exception:  # noqa: BLE001
                        pass
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:  # noqa: BLE001
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, n in names.items():
                        if r & bit:
                            self.reasons.add(n)
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True).stdout
                    a, b = out.strip().split(",")
                    self.samples.append(int(a))
                    self.max_mhz = int(b)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.01)

    def __enter__(self):
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thread.join(timeout=2)

    def summary(self):
        snake = {"HwSlowdown": "hw_slowdown", "HwThermalSlowdown": "hw_thermal_slowdown", "SwThermalSlowdown":
                 "sw_thermal_slowdown", "SwPowerCap": "sw_power_cap", "HwPowerBrakeSlowdown": "hw_power_brake"}
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(snake.get(r, r) for r in self.reasons), "samples": len(self.samples),
                "power_w_max": round(max(self.power), 1) if self.power else None,
                "power_w_median": round(statistics.median(self.power), 1) if self.power else None}


def dist_setup(n_gpus: int):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG", "WARN")   # (INFO / VERSION print the NCCL banner on stdout, next to the JSON line)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def max_over_ranks(ms: float, world: int) -> float:
    if world == 1:
        return ms
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def time_steps(fn, steps: int, warmup: int, world: int):
    """W warm-ups, then EXACTLY `steps` calls bracketed by barrier+synchronize, CUDA events on the current stream."""
    import torch
    for _ in range(warmup):
        fn()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier(world)
    return max_over_ranks(ms, world) / steps


def cpu_baseline_prefill(seconds_budget: float = 12.0):
    """The oracle (fp32 PyTorch restatement of the reference tests' oracle) timed on the host cores.
    Primary figure: BASELINE.json configs[0] / BASELINE.md section 3 -- fp32, B=2, S=1024, H=8, D=128, causal (the
    reference's own CPU-runnable case).  Second key: a bounded sample of the headline workload itself (whole heads of
    one batch element at S=4096, D=128, bf16 inputs, non-causal)."""
    import torch
    from oracle import ref_attention as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(0)
    # ---- config 1
    B1, S1, H1, D1 = 2, 1024, 8, 128
    q1, k1, v1 = (torch.randn(B1, S1, H1, D1, generator=g) for _ in range(3))
    R.attention_ref(q1[:1, :, :1], k1[:1, :, :1], v1[:1, :, :1], True)  # warm-up
    t0 = time.perf_counter()
    reps1 = 0
    while True:
        R.attention_ref(q1, k1, v1, True)
        reps1 += 1
        if time.perf_counter() - t0 > seconds_budget or reps1 >= 20:
            break
    dt1 = (time.perf_counter() - t0) / reps1
    fl1 = 4.0 * B1 * H1 * D1 * S1 * (S1 + 1) / 2
    # ---- bounded sample of the headline workload
    S, D = WORKLOAD["S"], WORKLOAD["D"]
    heads = 4
    q = torch.randn(1, S, heads, D, generator=g).to(torch.bfloat16)
    k = torch.randn(1, S, heads, D, generator=g).to(torch.bfloat16)
    v = torch.randn(1, S, heads, D, generator=g).to(torch.bfloat16)
    t0 = time.perf_counter()
    reps = 0
    while True:
        R.attention_ref(q, k, v)
        reps += 1
        if time.perf_counter() - t0 > seconds_budget or reps >= 6:
            break
    dt = (time.perf_counter() - t0) / reps
    flops = 4.0 * heads * S * S * D
    return {"value": round(fl1 / dt1 / 1e12, 4), "unit": "TFLOP/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"BASELINE config 1: fp32 PyTorch oracle, B={B1} S={S1} H={H1} D={D1} causal, whole problem, {reps1} reps, "
                      f"{dt1 * 1e3:.0f} ms each (causal flops = 4*B*H*D*S*(S+1)/2)",
            "headline_sample": {"value": round(flops / dt / 1e12, 4), "unit": "TFLOP/s",
                                "sample": f"{heads} of the workload's {WORKLOAD['B'] * WORKLOAD['H']} (batch, head) slices at "
                                          f"S={S}, D={D}, non-causal, {reps} reps, {dt * 1e3:.0f} ms each"}}


def load_reference_ext():
    import importlib.util
    so = os.path.join(ROOT, "oracle", "_ref", "mini_flash_attention_ref_C.so")
    if not os.path.exists(so):
        return None
    try:
        spec = importlib.util.spec_from_file_location("mini_flash_attention_ref_C", so)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception as e:  # noqa: BLE001
        sys.stderr.write(f"[bench] reference extension failed to load: {e}\n")
        return None


def profile_traffic(kernel_key: str):
    """(dram bytes per launch, source) from the committed ncu summary -- a stored capture, NOT measured in this run."""
    p = os.path.join(ROOT, "profiles", "ncu_summary.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p)).get(kernel_key, {})
            return d.get("dram_bytes_per_launch"), f"profiles/ncu_summary.json ({d.get('tag', 'stored ncu --set full capture')}); not re-measured in this run"
        except Exception:  # noqa: BLE001
            return None, None
    return None, None


def bind_to_gpu_numa_node(local: int):
    """Pin this process (and so its pinned-host allocations, first touch) to the CPUs next to its GPU.  Returns a
    description for the bench line; never fails the run."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        node = -1
        for cand in (bus, "0000:" + bus.split(":", 1)[-1]):
            f = f"/sys/bus/pci/devices/{cand}/numa_node"
            if os.path.exists(f):
                node = int(open(f).read().strip())
                break
        if node < 0:
            return {"numa_node": node, "bound": False, "why": "no NUMA node reported for the GPU"}
        cl = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        cpus = set()
        for part in cl.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if not use:
            return {"numa_node": node, "bound": False, "why": "none of the node's CPUs is in this process' cpuset"}
        os.sched_setaffinity(0, use)
        return {"numa_node": node, "bound": True, "cpus": len(use)}
    except Exception as e:  # noqa: BLE001
        return {"bound": False, "why": f"{type(e).__name__}: {e}"[:120]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-long-context", action="store_true")
    ap.add_argument("--e2e-sequential", action="store_true", help="ours: time the un-pipelined copy-kernel-copy sequence")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 2 s back-to-back leg")
    ap.add_argument("--no-libs", action="store_true", help="skip the flash-attn comparison row")
    ap.add_argument("--sustained-seconds", type=float, default=2.5)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    import torch
    if args.impl == "reference":
        # the reference arm runs on rank 0 alone (no process group); other ranks exit without work
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        rank, world, local = 0, 1, int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
    else:
        rank, world, local = dist_setup(args.gpus)
    n_gpus = world
    numa = bind_to_gpu_numa_node(local)
    pk = peaks()
    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float16
    B, H, S, D = (WORKLOAD[x] for x in "BHSD")
    flops = 4.0 * B * H * S * S * D

    ref_ext = None
    if args.impl == "ours":
        from mini_flash_attention_b200 import _lib, build
        build.build()
        import mini_flash_attention as mfa
        fwd = lambda q, k, v: mfa.flash_attn_func(q, k, v, causal=False)  # noqa: E731
        launch_count = _lib.launch_count
    else:
        ref_ext = load_reference_ext()
        if ref_ext is None:
            cb = cpu_baseline_prefill()
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "TFLOP/s",
                              "n_gpus": 1, "steps": 1, "warmup": 1, "higher_is_better": True, "scaling": "weak",
                              "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                              "config": {"workload": "prefill hd128 B48 H24 S4096 non-causal (bounded sample)"},
                              "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "TFLOP/s",
                                                          "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
            return 0
        fwd = lambda q, k, v: ref_ext.mini_flash_attention_forward(q, k, v, None, False, -1, -1)  # noqa: E731
        launch_count = lambda: 0  # noqa: E731

    g = torch.Generator(device="cuda").manual_seed(rank)
    q = torch.randn(B, S, H, D, device="cuda", dtype=dt, generator=g)
    k = torch.randn(B, S, H, D, device="cuda", dtype=dt, generator=g)
    v = torch.randn(B, S, H, D, device="cuda", dtype=dt, generator=g)

    with torch.inference_mode():
        # ---- device-resident leg
        with ClockSampler(local) as cs:
            l0 = launch_count()
            ms = time_steps(lambda: fwd(q, k, v), args.steps, args.warmup, world)
            launches = (launch_count() - l0 - args.warmup) if args.impl == "ours" else args.steps  # reference: one kernel per step
        clocks = cs.summary()
        value = n_gpus * flops / (ms * 1e-3) / 1e12

        # ---- sustained leg: the same launch back to back for >= 2 s with clocks and power sampled.  The headline's 20
        # launches (~0.2 s) sit between the burst and the sustained regime of the board; this is the steady state, and
        # its honest denominator is the SUSTAINED matmul figure of MEASURED_PEAKS.json.
        sustained = None
        if not args.no_sustained:
            n_sus = max(40, int(args.sustained_seconds * 1e3 / ms) + 1)
            with ClockSampler(local) as cs2:
                ms_sus = time_steps(lambda: fwd(q, k, v), n_sus, 3, world)
            tf_sus = flops / (ms_sus * 1e-3) / 1e12
            sustained = {"launches": n_sus, "seconds": round(n_sus * ms_sus * 1e-3, 2), "ms_per_step": round(ms_sus, 4),
                         "value": round(n_gpus * tf_sus, 2), "unit": "TFLOP/s", "clocks": cs2.summary(),
                         "frac_of_sustained_peak": round(tf_sus / pk["tflops_sustained"], 4) if pk.get("tflops_sustained") else None,
                         "frac_of_burst_peak": round(tf_sus / pk["tflops"], 4)}

        # ---- library row on the same box (comparison only): the official flash-attn wheel on the same tensors, the
        # bar the reference's own benchmark uses (benchmark/compare_prefill.py:42-55)
        libs = None
        if args.impl == "ours" and not args.no_libs and rank == 0:
            try:
                import flash_attn
                from flash_attn import flash_attn_func as _fa
                fa_ms = time_steps(lambda: _fa(q, k, v, causal=False), 5, 3, 1)
                libs = {"flash_attn": {"version": flash_attn.__version__, "ms_per_step": round(fa_ms, 3),
                                       "value": round(flops / (fa_ms * 1e-3) / 1e12, 1), "unit": "TFLOP/s",
                                       "max_abs_diff_vs_ours": float((_fa(q[:1], k[:1], v[:1]).float() - fwd(q[:1], k[:1], v[:1]).float()).abs().max())}}
            except Exception as e:  # noqa: BLE001
                libs = {"flash_attn": {"unavailable": f"{type(e).__name__}: {e}"[:160]}}
        if world > 1:
            barrier(world)

        # ---- end-to-end leg: pinned host buffers, H2D + kernel + D2H every step
        e2e = None
        head_sharded = None
        if not args.no_e2e:
            hq, hk, hv = (t.cpu().pin_memory() for t in (q, k, v))
            ho = torch.empty(q.shape, dtype=dt).pin_memory()

            if args.impl == "ours" and not args.e2e_sequential:
                # the repo's host-buffer entry point: H2D / kernel / D2H pipelined over batch chunks on three streams
                from mini_flash_attention_b200.host import flash_attn_func_host
                e2e_method = "flash_attn_func_host: H2D | kernel | D2H pipelined over 16 batch chunks, 3 streams"

                def step_e2e():
                    flash_attn_func_host(hq, hk, hv, False, out=ho)
            else:
                # what a caller of the reference does with host data: three copies in, the kernel, one copy out
                e2e_method = "sequential: q,k,v .to(cuda), one kernel, out copied back"

                def step_e2e():
                    dq = hq.to("cuda", non_blocking=True)
                    dk = hk.to("cuda", non_blocking=True)
                    dv = hv.to("cuda", non_blocking=True)
                    ho.copy_(fwd(dq, dk, dv), non_blocking=True)

            e2e_steps = max(3, min(args.steps, 10))
            ms_e2e = time_steps(step_e2e, e2e_steps, 3, world)
            nbytes = q.numel() * q.element_size()
            e2e = {"value": round(n_gpus * flops / (ms_e2e * 1e-3) / 1e12, 3), "unit": "TFLOP/s",
                   "h2d_bytes_per_step": 3 * nbytes, "d2h_bytes_per_step": nbytes, "ms_per_step": round(ms_e2e, 3),
                   "steps": e2e_steps, "method": e2e_method, "numa": numa,
                   "pcie_h2d_GBps_per_rank": round(3 * nbytes / (ms_e2e * 1e-3) / 1e9, 1),
                   "pcie_d2h_GBps_per_rank": round(nbytes / (ms_e2e * 1e-3) / 1e9, 1)}
            del hq, hk, hv, ho

        # ---- decode leg of the metric (ours only; GB/s of algorithmic KV bytes)
        decode = None
        if args.impl == "ours" or ref_ext is not None:
            del q, k, v
            torch.cuda.empty_cache()
            dB, dH, dD, KV, page = (DECODE[x] for x in ("B", "H", "D", "KV", "page"))
            per = KV // page
            nb = dB * per
            dq = torch.randn(dB, 1, dH, dD, device="cuda", dtype=dt, generator=g)
            kc = torch.randn(nb, page, dH, dD, device="cuda", dtype=dt, generator=g)
            vc = torch.randn(nb, page, dH, dD, device="cuda", dtype=dt, generator=g)
            bt = torch.randperm(nb, device="cuda", generator=g).reshape(dB, per).int()
            sl = torch.full((dB,), KV, dtype=torch.int32, device="cuda")
            if args.impl == "ours":
                dfn = lambda: mfa.flash_attn_with_kvcache(dq, kc, vc, sl, bt, 0)  # noqa: E731
            else:
                dfn = lambda: ref_ext.mini_flash_attention_with_kvcache(dq, kc, vc, sl, bt, False, 0)  # noqa: E731
            dms = time_steps(dfn, max(args.steps, 20), max(args.warmup, 5), world)
            dbytes = 2.0 * dB * dH * KV * dD * 2 + 2.0 * dB * dH * dD * 2  # K+V once per KV head, + q and o
            gbs = dbytes / (dms * 1e-3) / 1e9
            decode = {"workload": "decode B24 H24 D128 KV8192 paged(256) num_splits=auto", "ms_per_step": round(dms, 4),
                      "value": round(n_gpus * gbs, 1), "unit": "GB/s",
                      "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": pk["hbm"], "unit": "GB/s",
                                   "frac": round(gbs / pk["hbm"], 4), "traffic": profile_traffic("decode")[0],
                                   "traffic_source": profile_traffic("decode")[1]}}

        # ---- long-context decode, KV axis sharded over the ranks (BASELINE.json configs[4]; ours only): 1M rows in total,
        # exchange fused into the merge kernel over NVLink peer memory (no NCCL call on the data path).  STRONG scaling:
        # the total work is fixed, so `value` is the aggregate GB/s of the whole 1M-row problem.
        long_ctx = None
        if args.impl == "ours" and not args.no_long_context:
            try:
                from mini_flash_attention_b200 import parallel as par
                del dq, kc, vc, bt, sl
                torch.cuda.empty_cache()
                lB, lHq, lHkv, lKV, lpage = 1, 32, 8, 1 << 20, 256
                rows = par.kv_shard_rows(lKV, world, align=lpage)
                lnb = lB * rows // lpage
                lq = torch.randn(lB, 1, lHq, dD, device="cuda", dtype=dt, generator=torch.Generator(device="cuda").manual_seed(7))
                lk = torch.randn(lnb, lpage, lHkv, dD, device="cuda", dtype=dt)
                lv = torch.randn(lnb, lpage, lHkv, dD, device="cuda", dtype=dt)
                lbt = torch.randperm(lnb, device="cuda").reshape(lB, rows // lpage).int()
                lsl = par.kv_shard_lengths(torch.full((lB,), lKV, dtype=torch.int32, device="cuda"), rank, rows)
                xchg = par.KVShardExchange(world, rank, lB * lHq, dD)
                lfn = lambda: par.decode_kv_sharded(lq, lk, lv, lsl, lbt, world=world, exchange=xchg)  # noqa: E731
                # (1) parity of the N-GPU path, driver-visible: a 64K-row problem generated identically on every rank,
                # sharded over the ranks through the same fused exchange, against the UNSHARDED single-GPU call
                pKV = 65536
                prow = par.kv_shard_rows(pKV, world, align=lpage)
                pg = torch.Generator(device="cuda").manual_seed(99)
                pk_full = torch.randn(pKV // lpage, lpage, lHkv, dD, device="cuda", dtype=dt, generator=pg)
                pv_full = torch.randn(pKV // lpage, lpage, lHkv, dD, device="cuda", dtype=dt, generator=pg)
                pbt = torch.arange(pKV // lpage, device="cuda", dtype=torch.int32).reshape(1, -1)
                plen = torch.full((1,), pKV - 77, dtype=torch.int32, device="cuda")
                p_un = mfa.flash_attn_with_kvcache(lq, pk_full, pv_full, plen, pbt, 0)
                p_sh = par.decode_kv_sharded(lq, pk_full, pv_full, par.kv_shard_lengths(plen, rank, prow),
                                             par.kv_shard_block_table(pbt, rank, prow, lpage), world=world, exchange=xchg)
                perr = (p_sh.float() - p_un.float()).abs().max().reshape(1).double()
                if world > 1:
                    import torch.distributed as dist
                    dist.all_reduce(perr, op=dist.ReduceOp.MAX)
                del pk_full, pv_full
                # (2) eager steps (host call every step) and (3) the same step captured once in a CUDA graph and replayed:
                # the launch parameters never change (the exchange's step counter lives in device memory)
                lms_eager = time_steps(lfn, max(args.steps, 30), max(args.warmup, 5), world)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    for _ in range(3):
                        lfn()
                torch.cuda.current_stream().wait_stream(side)
                barrier(world)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    lout = lfn()
                lms = time_steps(graph.replay, max(args.steps, 30), max(args.warmup, 5), world)
                st = xchg.status()
                lbytes = 2.0 * lB * lKV * lHkv * dD * 2
                lgbs = lbytes / (lms * 1e-3) / 1e9
                long_ctx = {"workload": f"decode B1 Hq32 Hkv8 D128 KV=1M rows sharded over {world} GPU(s), paged(256)",
                            "scaling": "strong", "ms_per_step": round(lms, 4), "value": round(lgbs, 1), "unit": "GB/s",
                            "per_gpu": round(lgbs / world, 1), "frac_of_hbm_peak_per_gpu": round(lgbs / world / pk["hbm"], 4),
                            "driven_by": "CUDA graph replay of one captured step", "ms_per_step_eager": round(lms_eager, 4),
                            "exchange": "fused into the merge kernel: P2P push + flag over NVLink, no NCCL call",
                            "exchange_timeout_status": st, "launches_per_step": 2,
                            "parity": {"what": f"{world}-GPU KV-sharded result vs the unsharded single-GPU call, 64K-row cache, "
                                               "max over ranks", "max_abs_err": float(perr.item()),
                                       "tol": 1e-2 if args.dtype == "bf16" else 2e-3,
                                       "ok": bool(float(perr.item()) <= (1e-2 if args.dtype == "bf16" else 2e-3))}}
                del graph, lout
                xchg.close()
                del lk, lv
            except Exception as e:  # the headline legs above stay valid
                long_ctx = {"error": f"{type(e).__name__}: {e}"[:300]}

        # ---- head-sharded prefill (BASELINE.json configs[4]): ONE sequence of S=16K, Hq=32, Hkv=8, causal, its KV-head
        # groups cut over the ranks (parallel.prefill_sharded, no collective).  STRONG scaling: total work fixed.
        head_sharded = None
        if args.impl == "ours" and not args.no_long_context:
            try:
                from mini_flash_attention_b200 import parallel as par
                torch.cuda.empty_cache()
                hS, hHq, hHkv = 16384, 32, 8
                hg = torch.Generator(device="cuda").manual_seed(5)
                hq_ = torch.randn(1, hS, hHq, D, device="cuda", dtype=dt, generator=hg)
                hk_ = torch.randn(1, hS, hHkv, D, device="cuda", dtype=dt, generator=hg)
                hv_ = torch.randn(1, hS, hHkv, D, device="cuda", dtype=dt, generator=hg)
                res = {}
                for causal in (False, True):
                    hms = time_steps(lambda: par.prefill_sharded(hq_, hk_, hv_, causal, rank=rank, world=world), 10, 3, world)
                    hfl = 4.0 * hHq * D * (hS * (hS + 1) / 2 if causal else float(hS) * hS)
                    res["causal" if causal else "non_causal"] = {"ms_per_step": round(hms, 4), "value": round(hfl / (hms * 1e-3) / 1e12, 1),
                                                                 "unit": "TFLOP/s"}
                # parity of this rank's shard against the unsharded call on the same tensors (max over ranks)
                o_sh, shard = par.prefill_sharded(hq_, hk_, hv_, True, rank=rank, world=world)
                herr = torch.zeros(1, device="cuda", dtype=torch.float64)
                if o_sh is not None:
                    h_lo, h_hi = shard.q_heads(hHq // hHkv)
                    full = mfa.flash_attn_func(hq_, hk_, hv_, causal=True)
                    herr = (o_sh.float() - full[shard.batch_lo:shard.batch_hi, :, h_lo:h_hi].float()).abs().max().reshape(1).double()
                if world > 1:
                    import torch.distributed as dist
                    dist.all_reduce(herr, op=dist.ReduceOp.MAX)
                head_sharded = {"workload": f"prefill B1 S16384 Hq32 Hkv8 D128, KV-head groups over {world} GPU(s), no collective",
                                "scaling": "strong", **res, "parity_max_abs_err_vs_unsharded": float(herr.item())}
                del hq_, hk_, hv_
            except Exception as e:  # noqa: BLE001
                head_sharded = {"error": f"{type(e).__name__}: {e}"[:300]}

    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    out = {
        "metric": METRIC, "value": round(value, 2), "unit": "TFLOP/s", "n_gpus": n_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": "prefill flash_attn_func hd128 B48 H24 S4096 non-causal", "per_gpu_batch": B,
                   "l2": "inputs+output 4.8 GB per GPU >> 126 MB L2, no flush needed",
                   "parallelism": f"batch-sharded x{n_gpus}, no collective"},
        "clocks": clocks, "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": round(value / n_gpus, 2), "peak": pk["tflops"], "unit": "TFLOP/s",
                     "frac": round(value / n_gpus / pk["tflops"], 4), "traffic": profile_traffic("prefill")[0],
                     "traffic_source": profile_traffic("prefill")[1],
                     "peak_source": pk["source"] + ", burst figure (20 x ~8 ms launches)",
                     "frac_of_nominal_2250": round(value / n_gpus / 2250.0, 4)},
    }
    if sustained is not None:
        out["sustained"] = sustained
    if libs is not None:
        out["libs"] = libs
    if head_sharded is not None:
        out["head_sharded_prefill"] = head_sharded
    if e2e is not None:
        out["e2e"] = e2e
    if decode is not None:
        out["decode"] = decode
    if long_ctx is not None:
        out["long_context_decode"] = long_ctx
    if args.impl == "reference":
        out["impl"] = "reference"
        out["cpu_baseline"] = {"value": out["value"], "unit": "TFLOP/s", "cores": 0, "kind": "reference",
                               "sample": "the reference has no CPU implementation (csrc/mfa/api.cpp:123 CUDAGuard); "
                                         "this arm runs its unmodified CUDA kernels (oracle/_ref) on the same GPU"}
        if e2e is None:
            out["e2e"] = {"value": out["value"], "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    elif not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_prefill()
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
