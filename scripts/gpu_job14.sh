#!/bin/bash
# round-2 GPU pass M: TMEM load split is in; A/B of the polynomial share of the exponentials (FMA pipe instead of MUFU)
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 600 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py -q -x > $O/g14_pytest.log 2>&1; tail -3 $O/g14_pytest.log
MFA_B200_SHIM=py MFA_B200_LIB=$V/libpoly4.so timeout 600 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py -q -x > $O/g14_pytest_poly4.log 2>&1; tail -3 $O/g14_pytest_poly4.log
LIBS="default $V/libpoly8.so $V/libpoly6.so $V/libpoly4.so $V/libpoly3.so $V/libpoly2.so default $V/libpoly4.so"
timeout 900 python scripts/tune_prefill.py $LIBS > $O/g14_tune.txt 2>&1
echo "--- causal" >> $O/g14_tune.txt
TUNE_CAUSAL=1 timeout 900 python scripts/tune_prefill.py $LIBS >> $O/g14_tune.txt 2>&1
echo "--- S=1024" >> $O/g14_tune.txt
TUNE_S=1024 timeout 600 python scripts/tune_prefill.py default $V/libpoly4.so $V/libpoly8.so >> $O/g14_tune.txt 2>&1
grep -v Traceback $O/g14_tune.txt | cut -c1-250
for T in trace trace_poly4; do
MFA_B200_SHIM=py MFA_B200_LIB=$V/lib$T.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g14_$T.txt 2>&1; tail -5 $O/g14_$T.txt | cut -c1-250
done
