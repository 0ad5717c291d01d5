#!/bin/bash
# item-walk mode as a template parameter (fixed-length kernel = the plain strided walk again): suite + A/B
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1500 python -m pytest tests -q -x -m gpu > $O/g38_pytest.log 2>&1; tail -2 $O/g38_pytest.log
timeout 300 python scripts/varlen_ab.py > $O/g38_varlen.txt 2>&1; cat $O/g38_varlen.txt
for S in 512 4096 1024; do
echo "--- S=$S" >> $O/g38_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libhead.so >> $O/g38_tune.txt 2>&1
done
echo "--- S=4096 causal" >> $O/g38_tune.txt
TUNE_S=4096 TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libhead.so >> $O/g38_tune.txt 2>&1
grep -o '"lib": "[a-z.]*"\|"tflops_b2b": \[[0-9., ]*\]\|^---.*' $O/g38_tune.txt | paste - - - - - | head
