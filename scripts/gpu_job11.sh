#!/bin/bash
# round-2 GPU pass J (8 GPUs): torchrun parity at world 2/4/8, bench at N = 8, 4, 2 (ours) and the reference arm at N = 8
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_multigpu.py tests/test_kv_exchange_gpu.py -q -s > $O/g11_pytest_multigpu.log 2>&1; tail -3 $O/g11_pytest_multigpu.log
: > $O/scale_r02a.jsonl
for N in 8 4 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540+N)) bench.py --gpus $N --no-cpu-baseline > $O/g11_bench_n$N.json 2> $O/g11_bench_n$N.err
  grep "^{" $O/g11_bench_n$N.json >> $O/scale_r02a.jsonl; tail -2 $O/g11_bench_n$N.err | cut -c1-300
done
timeout 600 python bench.py --no-cpu-baseline >> $O/scale_r02a.jsonl 2> $O/g11_bench_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 8 --impl reference > $O/g11_bench_ref_n8.json 2> $O/g11_bench_ref_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29556 scripts/multi_gpu_check.py > $O/g11_multigpu_check.jsonl 2> $O/g11_multigpu_check.err; tail -3 $O/g11_multigpu_check.jsonl | cut -c1-400
python - <<'PY'
import json
for l in open('gpurun_out/scale_r02a.jsonl'):
    d=json.loads(l)
    lc=d.get('long_context_decode',{}); hs=d.get('head_sharded_prefill',{})
    print(d['n_gpus'], d['value'], d['ms_per_step'], 'decode', d['decode']['value'], 'long', lc.get('ms_per_step'), lc.get('ms_per_step_eager'), lc.get('parity'), 'hs', hs.get('non_causal'), hs.get('causal'), hs.get('parity_max_abs_err_vs_unsharded'), 'e2e', d['e2e']['value'], d['e2e'].get('pcie_h2d_GBps_per_rank'))
PY
