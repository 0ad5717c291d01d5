#!/bin/bash
# round-2 GPU pass I: early q_empty A/B, exchange kernel with one fence, parity of the touched paths
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_append_bottomright_gpu.py tests/test_regressions_gpu.py tests/test_kv_exchange_gpu.py tests/test_live_reference_gpu.py -q > $O/g10_pytest.log 2>&1; tail -2 $O/g10_pytest.log
LIBS="default $V/liblateq.so default $V/liblateq.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g10_tune.txt 2>&1
echo "--- causal" >> $O/g10_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g10_tune.txt 2>&1
for S in 2048 1024 512 256; do
echo "--- S=$S" >> $O/g10_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/liblateq.so >> $O/g10_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g10_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/liblateq.so >> $O/g10_tune.txt 2>&1
done
cat $O/g10_tune.txt
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-libs > $O/g10_bench.json 2> $O/g10_bench.err; cat $O/g10_bench.json
python scripts/profile_cases.py xchg 2 > $O/plain_case.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:xchg_merge --csv --log-file $O/g10_xchg_time.csv python scripts/profile_cases.py xchg 2 > $O/ncu_xchg.log 2>&1; grep xchg $O/g10_xchg_time.csv | tail -2
