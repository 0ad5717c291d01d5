#!/bin/bash
# what the driver runs at round end, on one GPU: smoke, the GPU test suite, both bench arms
O=gpurun_out; mkdir -p $O/profiles
python -c "import __graft_entry__ as g; g.smoke()" > $O/final_smoke.txt 2>&1; tail -1 $O/final_smoke.txt
timeout 1500 python -m pytest tests -x -q -m gpu > $O/final_pytest.log 2>&1; tail -2 $O/final_pytest.log
timeout 600 python bench.py --impl reference > $O/profiles/bench_ref_r02g.json 2> $O/final_bench_ref.err; cut -c1-300 $O/profiles/bench_ref_r02g.json
timeout 900 python bench.py > $O/profiles/bench_ours_r02g.json 2> $O/final_bench.err; cut -c1-300 $O/profiles/bench_ours_r02g.json
timeout 600 python scripts/sweep.py > $O/profiles/sweep_r02g.jsonl 2> $O/final_sweep.err; grep varlen $O/profiles/sweep_r02g.jsonl | cut -c1-200
