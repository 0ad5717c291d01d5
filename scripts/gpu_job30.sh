#!/bin/bash
# round-2 GPU pass Z: S MMA split into key halves, the hi half issued a P.V batch early -- parity under the variant, A/B, trace
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
MFA_B200_SHIM=py MFA_B200_LIB=$V/libssplit.so timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py -q -x -m gpu > $O/g30_pytest.log 2>&1; tail -2 $O/g30_pytest.log
LIBS="default $V/libssplit.so default $V/libssplit.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g30_tune.txt 2>&1
echo "--- causal" >> $O/g30_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py default $V/libssplit.so >> $O/g30_tune.txt 2>&1
for S in 1024 512; do
echo "--- S=$S" >> $O/g30_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libssplit.so >> $O/g30_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g30_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libssplit.so >> $O/g30_tune.txt 2>&1
done
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace_ssplit.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g30_trace.txt 2>&1; tail -6 $O/g30_trace.txt | cut -c1-250
