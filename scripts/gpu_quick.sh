#!/bin/bash
# quick GPU iteration: parity tests then the device-resident bench legs
TAG=${1:-q}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee $O/pytest_$TAG.log
python bench.py --no-cpu-baseline --no-e2e --steps 20 --warmup 5 2>$O/bench_$TAG.err | tee $O/bench_$TAG.json
tail -3 $O/bench_$TAG.err
