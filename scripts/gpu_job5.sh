#!/bin/bash
# round-2 GPU pass D: full parity suite, prefill speculative-exp A/B, persistent decode A/B, traces, bench
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1500 python -m pytest tests -m gpu -q -s > $O/g5_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g5_pytest.log; grep -E "passed|failed|reference decode rows" $O/g5_pytest.log | tail -5
LIBS="default $V/libnospec.so default $V/libnospec.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g5_tune.txt 2>&1
echo "--- causal" >> $O/g5_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g5_tune.txt 2>&1
for S in 2048 1024 512 256; do
echo "--- S=$S" >> $O/g5_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libnospec.so >> $O/g5_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g5_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libnospec.so >> $O/g5_tune.txt 2>&1
done
cat $O/g5_tune.txt
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g5_trace.txt 2>&1
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace_nospec.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g5_trace_nospec.txt 2>&1
tail -5 $O/g5_trace.txt; tail -5 $O/g5_trace_nospec.txt
timeout 900 python scripts/decode_ab.py MFA_DECODE_NONPERSISTENT=1 > $O/g5_decode_ab.txt 2>&1; cat $O/g5_decode_ab.txt
timeout 900 python bench.py > $O/g5_bench.json 2> $O/g5_bench.err; cat $O/g5_bench.json; tail -3 $O/g5_bench.err
