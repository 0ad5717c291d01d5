#!/bin/bash
# DPAD = 192 instantiation of the prefill kernel: parity (all prefill-side tests) + A/B against padding to 256
O=gpurun_out; mkdir -p $O/profiles
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_lse_gpu.py tests/test_regressions_gpu.py tests/test_append_bottomright_gpu.py tests/test_flash_attn_oracle_gpu.py tests/test_live_reference_gpu.py -q -x -m gpu > $O/g33_pytest.log 2>&1; tail -3 $O/g33_pytest.log
timeout 300 python scripts/d192_ab.py > $O/profiles/prefill_d192_ab_r02d.txt 2>&1
MFA_PREFILL_NO_D192=1 timeout 300 python scripts/d192_ab.py >> $O/profiles/prefill_d192_ab_r02d.txt 2>&1
cat $O/profiles/prefill_d192_ab_r02d.txt
