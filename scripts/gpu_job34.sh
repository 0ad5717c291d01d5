#!/bin/bash
# packed batches persistent with a snake-order item walk: parity + A/B
O=gpurun_out; mkdir -p $O/profiles
timeout 900 python -m pytest tests/test_varlen_gpu.py tests/test_prefill_gpu.py tests/test_regressions_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_append_bottomright_gpu.py tests/test_live_reference_gpu.py tests/test_flash_attn_oracle_gpu.py -q -x -m gpu > $O/g34_pytest.log 2>&1; tail -2 $O/g34_pytest.log
timeout 300 python scripts/varlen_ab.py > $O/profiles/varlen_persistent_ab_r02e.txt 2>&1
MFA_PREFILL_VARLEN_NONPERSISTENT=1 timeout 300 python scripts/varlen_ab.py >> $O/profiles/varlen_persistent_ab_r02e.txt 2>&1
cat $O/profiles/varlen_persistent_ab_r02e.txt
