#!/bin/bash
# round-2 GPU pass F: pipe-rate micro-benchmark, PRMT-pack A/B, bench
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 120 scripts/pipe_rates > $O/g7_pipe_rates.txt 2>&1; cat $O/g7_pipe_rates.txt
timeout 600 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_live_reference_gpu.py -q > $O/g7_pytest_default.log 2>&1; tail -2 $O/g7_pytest_default.log
MFA_B200_SHIM=py MFA_B200_LIB=$V/libprmt.so timeout 600 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_live_reference_gpu.py tests/test_flash_attn_oracle_gpu.py -q > $O/g7_pytest_prmt.log 2>&1; tail -2 $O/g7_pytest_prmt.log
LIBS="default $V/libprmt.so default $V/libprmt.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g7_tune.txt 2>&1
echo "--- causal" >> $O/g7_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g7_tune.txt 2>&1
for S in 2048 1024 512; do
echo "--- S=$S" >> $O/g7_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libprmt.so >> $O/g7_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g7_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libprmt.so >> $O/g7_tune.txt 2>&1
done
cat $O/g7_tune.txt
MFA_B200_LIB=$V/libprmt.so timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-long-context > $O/g7_bench_prmt.json 2> $O/g7_bench_prmt.err; cat $O/g7_bench_prmt.json
