#!/bin/bash
# round-2 GPU pass U: next item's coordinates pinned early in the MMA and softmax roles -- parity, A/B, trace
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py -q -x -m gpu > $O/g23_pytest.log 2>&1; tail -3 $O/g23_pytest.log
LIBS="default $V/libprev.so"
for S in 4096 1024 512 256; do
echo "--- S=$S" >> $O/g23_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g23_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g23_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 400 python scripts/tune_prefill.py $LIBS >> $O/g23_tune.txt 2>&1
done
grep -v Traceback $O/g23_tune.txt | cut -c1-250
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 512 > $O/g23_trace_s512.txt 2>&1; tail -17 $O/g23_trace_s512.txt | cut -c1-260
