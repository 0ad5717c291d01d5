#!/bin/bash
# round-2 GPU pass T: causal mask applied per 32-column chunk (chunks every row of the warp sees whole are skipped)
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py tests/test_live_reference_gpu.py tests/test_flash_attn_oracle_gpu.py -q -x -m gpu > $O/g21_pytest.log 2>&1; tail -3 $O/g21_pytest.log
LIBS="default $V/liboldmask.so"
for S in 4096 2048 1024 512 256; do
echo "--- S=$S causal" >> $O/g21_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 400 python scripts/tune_prefill.py $LIBS >> $O/g21_tune.txt 2>&1
done
echo "--- S=4096" >> $O/g21_tune.txt
timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g21_tune.txt 2>&1
grep -v Traceback $O/g21_tune.txt | cut -c1-250
