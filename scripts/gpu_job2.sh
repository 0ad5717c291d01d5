#!/bin/bash
# round-2 GPU pass A: parity of the flat-stream prefill kernel, A/B against the per-tile ring, trace, library rows
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,power.limit,clocks.max.sm --format=csv > $O/g2_smi.txt
timeout 900 python -m pytest tests/test_prefill_gpu.py tests/test_varlen_gpu.py tests/test_window_gpu.py tests/test_append_bottomright_gpu.py tests/test_lse_gpu.py -x -q > $O/g2_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g2_pytest.log; tail -3 $O/g2_pytest.log
timeout 300 python scripts/tune_prefill.py default mini_flash_attention_b200/variants/libnoflat.so default mini_flash_attention_b200/variants/libnoflat.so > $O/g2_tune.txt 2>&1
TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default mini_flash_attention_b200/variants/libnoflat.so >> $O/g2_tune.txt 2>&1
TUNE_S=1024 timeout 300 python scripts/tune_prefill.py default mini_flash_attention_b200/variants/libnoflat.so >> $O/g2_tune.txt 2>&1
TUNE_S=1024 TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default mini_flash_attention_b200/variants/libnoflat.so >> $O/g2_tune.txt 2>&1
cat $O/g2_tune.txt
MFA_B200_SHIM=py MFA_B200_LIB=mini_flash_attention_b200/variants/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g2_trace.txt 2>&1
timeout 200 python scripts/umma_rate.py > $O/g2_umma.txt 2>&1
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-long-context > $O/g2_bench.json 2> $O/g2_bench.err; cat $O/g2_bench.json
timeout 600 python scripts/libs_row.py --quick > $O/g2_libs.txt 2>&1; grep -v watchdog $O/g2_libs.txt | tail -8
