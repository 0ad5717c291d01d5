#!/bin/bash
# decode producer: first page id fetched together with the sequence length -- parity + A/B on the short-KV shapes
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_fp8_kv_gpu.py tests/test_kv_exchange_gpu.py tests/test_window_gpu.py tests/test_append_bottomright_gpu.py tests/test_live_reference_gpu.py tests/test_zz_no_alloc.py -q -x -m gpu > $O/g31_pytest.log 2>&1; tail -2 $O/g31_pytest.log
for R in 1 2; do
echo "=== new" >> $O/g31_splits.txt
timeout 300 python scripts/decode_splits_gqa.py quick >> $O/g31_splits.txt 2>&1
echo "=== head" >> $O/g31_splits.txt
MFA_B200_SHIM=py MFA_B200_LIB=$V/libhead.so timeout 300 python scripts/decode_splits_gqa.py quick >> $O/g31_splits.txt 2>&1
done
cut -c1-110 $O/g31_splits.txt
