#!/bin/bash
# round-2 GPU pass B: full parity suite, prefill variant A/B (spill-free MMA warp), trace, full bench
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1500 python -m pytest tests -m gpu -q > $O/g3_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g3_pytest.log; tail -15 $O/g3_pytest.log
LIBS="default $V/libnoflat.so $V/libnosplit.so $V/libwarparr.so $V/libh80.so $V/libnosplit_noflat.so $V/libnosplit_warparr.so default $V/libnoflat.so"
timeout 600 python scripts/tune_prefill.py $LIBS > $O/g3_tune.txt 2>&1
echo "--- causal" >> $O/g3_tune.txt
TUNE_CAUSAL=1 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g3_tune.txt 2>&1
echo "--- S=1024" >> $O/g3_tune.txt
TUNE_S=1024 timeout 300 python scripts/tune_prefill.py default $V/libnoflat.so $V/libnosplit.so $V/libnosplit_noflat.so >> $O/g3_tune.txt 2>&1
echo "--- S=1024 causal" >> $O/g3_tune.txt
TUNE_S=1024 TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libnoflat.so $V/libnosplit.so $V/libnosplit_noflat.so >> $O/g3_tune.txt 2>&1
echo "--- poly" >> $O/g3_tune.txt
MFA_PREFILL_POLY=1 timeout 300 python scripts/tune_prefill.py default $V/libnosplit.so >> $O/g3_tune.txt 2>&1
cat $O/g3_tune.txt
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g3_trace.txt 2>&1
timeout 900 python bench.py > $O/g3_bench.json 2> $O/g3_bench.err; cat $O/g3_bench.json; tail -3 $O/g3_bench.err
