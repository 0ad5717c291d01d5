#!/bin/bash
# round-2 GPU pass C: full parity suite (fp8 decode, fixed live-reference tests), prefill A/B (prewait, early S), trace, bench
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1500 python -m pytest tests -m gpu -q > $O/g4_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g4_pytest.log; tail -12 $O/g4_pytest.log
LIBS="default $V/libnoearly.so $V/libnoflat.so default $V/libnoearly.so $V/libnoflat.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g4_tune.txt 2>&1
echo "--- causal" >> $O/g4_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g4_tune.txt 2>&1
for S in 2048 1024 512 256; do
echo "--- S=$S" >> $O/g4_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libnoearly.so $V/libnoflat.so >> $O/g4_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g4_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libnoearly.so $V/libnoflat.so >> $O/g4_tune.txt 2>&1
done
cat $O/g4_tune.txt
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g4_trace.txt 2>&1
timeout 900 python bench.py > $O/g4_bench.json 2> $O/g4_bench.err; cat $O/g4_bench.json; tail -3 $O/g4_bench.err
