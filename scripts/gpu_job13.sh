#!/bin/bash
# round-2 GPU pass L: one thread per row again; A/B of the softmax-chain variants (TMEM load split, late publication of
# the first P half, P released in three parts), traces of the winners
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 600 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_decode_gpu.py -q -x > $O/g13_pytest.log 2>&1; tail -3 $O/g13_pytest.log
LIBS="default $V/libld.so $V/liblate8.so $V/libldlate8.so $V/libp3.so $V/libp3ld.so $V/libp3_8ld.so $V/liblate16.so default"
timeout 900 python scripts/tune_prefill.py $LIBS > $O/g13_tune.txt 2>&1
echo "--- causal" >> $O/g13_tune.txt
TUNE_CAUSAL=1 timeout 900 python scripts/tune_prefill.py $LIBS >> $O/g13_tune.txt 2>&1
echo "--- S=1024" >> $O/g13_tune.txt
TUNE_S=1024 timeout 600 python scripts/tune_prefill.py $LIBS >> $O/g13_tune.txt 2>&1
grep -v Traceback $O/g13_tune.txt | cut -c1-250
for T in trace trace_ldlate8 trace_p3ld; do
MFA_B200_SHIM=py MFA_B200_LIB=$V/lib$T.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g13_$T.txt 2>&1; tail -5 $O/g13_$T.txt | cut -c1-250
done
