#!/bin/bash
# round-2 GPU pass P: item-boundary work of the prefill kernel (staged epilogue with st.shared, Q tile 0 + first K ahead of
# Q tile 1, next item's descriptor computed under the waits) -- parity, A/B, traces (short non-causal and causal);
# decode: combine / exchange-merge as programmatic dependents of the split-KV kernel
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1200 python -m pytest tests -q -x -m gpu > $O/g17_pytest.log 2>&1; tail -3 $O/g17_pytest.log
LIBS="default $V/libqtogether.so $V/libepidirect.so"
timeout 600 python scripts/tune_prefill.py $LIBS default > $O/g17_tune.txt 2>&1
echo "--- causal" >> $O/g17_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g17_tune.txt 2>&1
for S in 2048 1024 512 256; do
echo "--- S=$S" >> $O/g17_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g17_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g17_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g17_tune.txt 2>&1
done
grep -v Traceback $O/g17_tune.txt | cut -c1-250
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 512 > $O/g17_trace_s512.txt 2>&1; tail -17 $O/g17_trace_s512.txt | cut -c1-260
for T in trace trace_epidirect; do
MFA_B200_SHIM=py MFA_B200_LIB=$V/lib$T.so timeout 200 python scripts/trace_prefill.py 1024 causal > $O/g17_${T}_s1024c.txt 2>&1; tail -16 $O/g17_${T}_s1024c.txt | cut -c1-300
done
timeout 600 python scripts/decode_splits_gqa.py small > $O/g17_splits_pdl.txt 2>&1; cat $O/g17_splits_pdl.txt
MFA_DECODE_NO_PDL=1 timeout 600 python scripts/decode_splits_gqa.py small > $O/g17_splits_nopdl.txt 2>&1; cat $O/g17_splits_nopdl.txt
