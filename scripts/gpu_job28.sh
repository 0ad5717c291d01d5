#!/bin/bash
# final state, one rank count per call: bash scripts/gpu_job28.sh <N>   (appends to gpurun_out/profiles/scale_r02c_n<N>.jsonl)
N=$1; O=gpurun_out; mkdir -p $O/profiles
if [ "$N" = "1" ]; then
  timeout 1200 python -m pytest tests -q -x -m gpu > $O/g28_pytest.log 2>&1; tail -2 $O/g28_pytest.log
  python -c "import __graft_entry__ as g; g.smoke()" > $O/g28_smoke.txt 2>&1; tail -1 $O/g28_smoke.txt
  timeout 900 python bench.py > $O/profiles/scale_r02c_n1.jsonl 2> $O/g28_bench_n1.err; cut -c1-400 $O/profiles/scale_r02c_n1.jsonl
  timeout 900 python scripts/sweep.py > $O/profiles/sweep_r02c.jsonl 2> $O/g28_sweep.err; grep -c . $O/profiles/sweep_r02c.jsonl
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540+N)) bench.py --gpus $N --no-cpu-baseline > $O/g28_bench_n$N.json 2> $O/g28_bench_n$N.err
  grep "^{" $O/g28_bench_n$N.json > $O/profiles/scale_r02c_n$N.jsonl; tail -2 $O/g28_bench_n$N.err | cut -c1-300; cut -c1-300 $O/profiles/scale_r02c_n$N.jsonl
fi
