#!/usr/bin/env python
"""Turn ncu artefacts from gpurun_out/ into the tracked summaries under profiles/.

    python scripts/ncu_summarize.py <tag> [--out DIR] [key ...]        e.g. r02a prefill decode decode_kv512

Reads  gpurun_out/launches_<tag>.csv, gpurun_out/<key>_<tag>.ncu-rep (needs `ncu` on PATH, no GPU; default keys
       prefill decode).  Runs on the GPU box too, so that only the text summaries travel back (--out gpurun_out/profiles).
Writes <out>/launches_<tag>.txt      per-kernel launch counts / times / share of the step
       <out>/<key>_<tag>.txt         key raw metrics + top stalled SASS instructions
       <out>/ncu_summary.json        {key: {...}} incl. dram_bytes_per_launch (bench.py reads prefill / decode)
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO = os.path.join(ROOT, "gpurun_out")
PR = os.path.join(ROOT, "profiles")

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.per_cycle_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.max",
    "sm__cycles_elapsed.max.per_second", "smsp__inst_executed.sum", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "smsp__sass_inst_executed_op_tmem_ldt.sum", "smsp__sass_inst_executed_op_tmem_stt.sum",
    "launch__shared_mem_per_block_dynamic",
]


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def launches(tag):
    p = os.path.join(GO, f"launches_{tag}.csv")
    if not os.path.exists(p):
        return
    lines = [l for l in open(p) if l.startswith('"')]
    rows = list(csv.reader(lines))
    ix = {h: i for i, h in enumerate(rows[0])}
    agg = {}
    for r in rows[1:]:
        name = r[ix["Kernel Name"]]
        short = name.split("(")[0].split("::")[-1][:70]
        if "mfa::" in name:
            short = "mfa::" + short
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        a = agg.setdefault(short, [0, 0.0])
        a[0] += 1
        a[1] += ns
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(PR, f"launches_{tag}.txt"), "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none, command: python bench.py --steps 2 --warmup 3 "
                f"--no-cpu-baseline --no-e2e  (tag {tag})\n# per-launch times are cold-cache and serialised: compare SHARES\n")
        f.write(f"{'kernel':72s} {'launches':>8s} {'total_us':>12s} {'avg_us':>10s} {'share':>7s}\n")
        for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k:72s} {n:8d} {ns / 1e3:12.1f} {ns / 1e3 / n:10.1f} {100 * ns / tot:6.1f}%\n")
    print(open(os.path.join(PR, f"launches_{tag}.txt")).read())


def kernel(tag, key):
    rep = os.path.join(GO, f"{key}_{tag}.ncu-rep")
    if not os.path.exists(rep):
        return None
    raw = ncu_csv(rep, "raw")
    if len(raw) < 3:
        return None
    hdr, units, vals = raw[0], raw[1], raw[2]
    m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    src = ncu_csv(rep, "source")
    shdr = src[1]
    ix = {h: i for i, h in enumerate(shdr)}
    body = [r for r in src[2:] if len(r) == len(shdr)]
    tot = sum(int(r[ix["# Samples"]] or 0) for r in body) or 1
    stall_cols = [h for h in shdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {c: sum(int(r[ix[c]] or 0) for r in body) for c in stall_cols}
    top = sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]] or 0))[:25]
    with open(os.path.join(PR, f"{key}_{tag}.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on, kernel {m.get('Kernel Name', ('', ''))[0][:120]}\n")
        for k in KEYS:
            if k in m:
                f.write(f"{k:75s} {m[k][0]:>18s} {m[k][1]}\n")
        f.write("\n# warp-stall samples by reason\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
            if v:
                f.write(f"{k:28s} {v:8d} {100 * v / tot:5.1f}%\n")
        f.write("\n# top SASS instructions by samples (index, samples, share, executed, SASS, top reasons)\n")
        for i in sorted(top):
            r = body[i]
            s = int(r[ix["# Samples"]] or 0)
            rs = sorted(((int(r[ix[c]] or 0), c) for c in stall_cols), reverse=True)[:2]
            f.write(f"{i:5d} {s:7d} {100 * s / tot:5.1f}% ex={r[ix['Instructions Executed']]:>10s}  {r[ix['Source']].strip()[:80]:80s} {rs}\n")

    def num(k):
        v, u = m.get(k, ("0", ""))
        x = float(v.replace(",", "") or 0)
        return x * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12, "ms": 1e-3, "us": 1e-6, "ns": 1e-9,
                    "s": 1}.get(u, 1)
    return {"tag": tag, "duration_s": num("gpu__time_duration.sum"),
            "dram_bytes_per_launch": num("dram__bytes_read.sum") + num("dram__bytes_write.sum"),
            "tensor_pipe_pct": num("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
            "xu_pipe_pct": num("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed"),
            "dram_pct": num("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            "registers": num("launch__registers_per_thread")}


def main():
    global PR
    args = sys.argv[1:]
    tag = args.pop(0)
    if "--out" in args:
        i = args.index("--out")
        PR = os.path.abspath(args[i + 1])
        del args[i:i + 2]
    keys = args or ["prefill", "decode"]
    os.makedirs(PR, exist_ok=True)
    launches(tag)
    sp = os.path.join(PR, "ncu_summary.json")
    summ = json.load(open(sp)) if os.path.exists(sp) else {}
    for key in keys:
        r = kernel(tag, key)
        if r:
            summ[key] = r
            print(key, r)
    json.dump(summ, open(sp, "w"), indent=1)


if __name__ == "__main__":
    main()
