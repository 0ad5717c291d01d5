#!/bin/bash
# round-2 GPU pass K: two softmax threads per row -- parity, A/B against one thread per row, trace, bench
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_append_bottomright_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_live_reference_gpu.py tests/test_flash_attn_oracle_gpu.py -q -x > $O/g12_pytest.log 2>&1; tail -3 $O/g12_pytest.log
LIBS="default $V/librs1.so default $V/librs1.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g12_tune.txt 2>&1
echo "--- causal" >> $O/g12_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g12_tune.txt 2>&1
for S in 2048 1024 512 256; do
echo "--- S=$S" >> $O/g12_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/librs1.so >> $O/g12_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g12_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/librs1.so >> $O/g12_tune.txt 2>&1
done
cat $O/g12_tune.txt
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g12_trace.txt 2>&1; tail -6 $O/g12_trace.txt | cut -c1-250
timeout 400 python bench.py --no-cpu-baseline --no-e2e --no-libs --no-long-context > $O/g12_bench.json 2> $O/g12_bench.err; cat $O/g12_bench.json; tail -2 $O/g12_bench.err
