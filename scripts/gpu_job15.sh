#!/bin/bash
# round-2 GPU pass N: split merge fused into the split-KV decode kernel (last-ticket CTA) -- parity, split sweep with and
# without the fusion; short-sequence prefill item timeline
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_fp8_kv_gpu.py tests/test_kv_exchange_gpu.py tests/test_shim.py tests/test_window_gpu.py tests/test_append_bottomright_gpu.py tests/test_lse_gpu.py tests/test_live_reference_gpu.py tests/test_flash_attn_oracle_gpu.py tests/test_regressions_gpu.py -q -x -m gpu > $O/g15_pytest.log 2>&1; tail -3 $O/g15_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/g15_smoke.txt 2>&1; tail -1 $O/g15_smoke.txt
timeout 900 python scripts/decode_splits_gqa.py > $O/g15_splits_fused.txt 2>&1; cat $O/g15_splits_fused.txt
MFA_DECODE_SEPARATE_COMBINE=1 timeout 600 python scripts/decode_splits_gqa.py quick > $O/g15_splits_separate.txt 2>&1; cat $O/g15_splits_separate.txt
for S in 512 1024; do
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py $S > $O/g15_trace_s$S.txt 2>&1; tail -22 $O/g15_trace_s$S.txt | cut -c1-260
done
