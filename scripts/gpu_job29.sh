#!/bin/bash
# round-2 GPU pass Y: epilogue reads the O row from TMEM with one wait per 128 columns -- parity + A/B
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py -q -x -m gpu > $O/g29_pytest.log 2>&1; tail -2 $O/g29_pytest.log
LIBS="default $V/libhead.so"
for S in 256 512 1024 4096; do
echo "--- S=$S" >> $O/g29_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g29_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g29_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g29_tune.txt 2>&1
done
