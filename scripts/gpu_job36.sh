#!/bin/bash
# ragged causal packed batches: device-built list of the items that exist, persistent snake walk -- parity + A/B
O=gpurun_out; mkdir -p $O/profiles
V=mini_flash_attention_b200/variants
timeout 1500 python -m pytest tests -q -x -m gpu > $O/g36_pytest.log 2>&1; tail -2 $O/g36_pytest.log
timeout 300 python scripts/varlen_ab.py > $O/profiles/varlen_compact_ab_r02e.txt 2>&1
MFA_PREFILL_NO_COMPACT=1 timeout 300 python scripts/varlen_ab.py >> $O/profiles/varlen_compact_ab_r02e.txt 2>&1
cat $O/profiles/varlen_compact_ab_r02e.txt
for S in 4096 512; do
echo "--- S=$S" >> $O/g36_tune.txt
TUNE_S=$S timeout 300 python scripts/tune_prefill.py default $V/libhead.so >> $O/g36_tune.txt 2>&1
echo "--- S=$S causal" >> $O/g36_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 300 python scripts/tune_prefill.py default $V/libhead.so >> $O/g36_tune.txt 2>&1
done
grep -o '"lib": "[a-z.]*"\|"tflops_b2b": \[[0-9., ]*\]\|^---.*' $O/g36_tune.txt | paste - - - | head -20
