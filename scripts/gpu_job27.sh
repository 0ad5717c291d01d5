#!/bin/bash
# round-2 final multi-GPU pass (8 GPUs): torchrun parity at world 2/4/8, bench at N = 8 and 2 (+ N = 1 for the scaling base),
# KV-sharded step with and without the dependent launch of the exchange-merge kernel
O=gpurun_out; mkdir -p $O $O/profiles
TAG=r02b
timeout 900 python -m pytest tests/test_multigpu.py tests/test_kv_exchange_gpu.py -q -s > $O/g27_pytest_multigpu.log 2>&1; tail -3 $O/g27_pytest_multigpu.log
: > $O/profiles/scale_$TAG.jsonl
for N in 8 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540+N)) bench.py --gpus $N --no-cpu-baseline > $O/g27_bench_n$N.json 2> $O/g27_bench_n$N.err
  grep "^{" $O/g27_bench_n$N.json >> $O/profiles/scale_$TAG.jsonl; tail -2 $O/g27_bench_n$N.err | cut -c1-300
done
timeout 600 python bench.py --no-cpu-baseline >> $O/profiles/scale_$TAG.jsonl 2> $O/g27_bench_n1.err
MFA_DECODE_NO_PDL=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29557 bench.py --gpus 8 --no-cpu-baseline --no-e2e --no-libs --no-sustained > $O/g27_bench_n8_nopdl.json 2> $O/g27_bench_n8_nopdl.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29556 scripts/multi_gpu_check.py > $O/profiles/multigpu_$TAG.jsonl 2> $O/g27_multigpu_check.err; tail -3 $O/profiles/multigpu_$TAG.jsonl | cut -c1-400
python - <<'PY'
import json
def show(d, tag=''):
    lc=d.get('long_context_decode',{}); hs=d.get('head_sharded_prefill',{})
    print(tag, d['n_gpus'], d['value'], d['ms_per_step'], 'decode', d['decode']['value'], 'long', lc.get('ms_per_step'), lc.get('ms_per_step_eager'), lc.get('parity'), 'hs', hs.get('non_causal'), hs.get('causal'), hs.get('parity_max_abs_err_vs_unsharded'), 'e2e', (d.get('e2e') or {}).get('value'))
for l in open('gpurun_out/profiles/scale_r02b.jsonl'): show(json.loads(l))
for l in open('gpurun_out/g27_bench_n8_nopdl.json'):
    if l.startswith('{'): show(json.loads(l), 'NO_PDL')
PY
