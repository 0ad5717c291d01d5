#!/bin/bash
# round-2 GPU pass R: epilogue form per tile (direct for a tile that ends before its item, staged otherwise) -- parity + A/B
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py tests/test_live_reference_gpu.py tests/test_flash_attn_oracle_gpu.py -q -x -m gpu > $O/g19_pytest.log 2>&1; tail -3 $O/g19_pytest.log
LIBS="default $V/liballstaged.so $V/libepidirect.so $V/libqsplit.so"
for S in 4096 1024 512 256; do
echo "--- S=$S causal" >> $O/g19_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 400 python scripts/tune_prefill.py $LIBS >> $O/g19_tune.txt 2>&1
done
for S in 4096 512; do
echo "--- S=$S" >> $O/g19_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g19_tune.txt 2>&1
done
grep -v Traceback $O/g19_tune.txt | cut -c1-250
