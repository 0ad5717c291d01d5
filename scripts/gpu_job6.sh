#!/bin/bash
# round-2 GPU pass E: parity suite, row-max variant A/B + fine trace, varlen order, fp8 eager, sweep + library rows
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1500 python -m pytest tests -m gpu -q -s > $O/g6_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g6_pytest.log; grep -E "passed|failed|reference decode rows|^FAILED" $O/g6_pytest.log | tail -8
LIBS="default $V/libmax2.so default $V/libmax2.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g6_tune.txt 2>&1
echo "--- causal" >> $O/g6_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g6_tune.txt 2>&1
for S in 1024; do
echo "--- S=$S" >> $O/g6_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libmax2.so >> $O/g6_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g6_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libmax2.so >> $O/g6_tune.txt 2>&1
done
cat $O/g6_tune.txt
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g6_trace.txt 2>&1
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace_max2.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g6_trace_max2.txt 2>&1
tail -5 $O/g6_trace.txt; tail -5 $O/g6_trace_max2.txt
rm -f $O/sweep.jsonl $O/libs.jsonl
timeout 1200 python scripts/sweep.py > $O/g6_sweep.txt 2>&1; tail -30 $O/g6_sweep.txt
timeout 1200 python scripts/libs_row.py > $O/g6_libs.txt 2>&1; grep -v watchdog $O/g6_libs.txt | tail -24
timeout 600 python scripts/decode_ab.py > $O/g6_decode_ab.txt 2>&1; grep e4m3 $O/g6_decode_ab.txt
