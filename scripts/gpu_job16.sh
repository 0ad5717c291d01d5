#!/bin/bash
# round-2 GPU pass O: smem-staged coalesced prefill epilogue vs direct row-per-thread stores; decode split heuristic v2,
# fused merge only for single-wave launches
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py tests/test_decode_gpu.py tests/test_live_reference_gpu.py -q -x -m gpu > $O/g16_pytest.log 2>&1; tail -3 $O/g16_pytest.log
LIBS="default $V/libepidirect.so default $V/libepidirect.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g16_tune.txt 2>&1
echo "--- causal" >> $O/g16_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py default $V/libepidirect.so >> $O/g16_tune.txt 2>&1
for S in 2048 1024 512 256; do
echo "--- S=$S" >> $O/g16_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libepidirect.so >> $O/g16_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g16_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libepidirect.so >> $O/g16_tune.txt 2>&1
done
grep -v Traceback $O/g16_tune.txt | cut -c1-250
for T in trace trace_epidirect; do
MFA_B200_SHIM=py MFA_B200_LIB=$V/lib$T.so timeout 200 python scripts/trace_prefill.py 512 > $O/g16_${T}_s512.txt 2>&1; tail -18 $O/g16_${T}_s512.txt | cut -c1-260
done
timeout 900 python scripts/decode_splits_gqa.py > $O/g16_splits.txt 2>&1; cat $O/g16_splits.txt
MFA_DECODE_SEPARATE_COMBINE=1 timeout 900 python scripts/decode_splits_gqa.py > $O/g16_splits_separate.txt 2>&1; cat $O/g16_splits_separate.txt
