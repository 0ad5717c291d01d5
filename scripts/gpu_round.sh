#!/bin/bash
# One GPU-box pass: parity tests, both bench arms, ncu launch list + full capture of the two hot kernels.
# Usage (from the repo root on the box):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_$TAG.log
tail -3 $O/pytest_gpu_$TAG.log
python bench.py --impl reference > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "bench ref rc=$?"
python bench.py > $O/bench_ours_$TAG.json 2> $O/bench_ours_$TAG.err; echo "bench ours rc=$?"
cat $O/bench_ref_$TAG.json $O/bench_ours_$TAG.json
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$SHORT > $O/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$TAG.csv $SHORT > $O/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
$SHORT > $O/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:prefill_fwd -s 3 -c 1 -f -o $O/prefill_$TAG $SHORT > $O/ncu_prefill_$TAG.log 2>&1
echo "ncu prefill rc=$?"
$SHORT > $O/plain3_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:decode_splitkv -s 5 -c 1 -f -o $O/decode_$TAG $SHORT > $O/ncu_decode_$TAG.log 2>&1
echo "ncu decode rc=$?"
ls -la $O
