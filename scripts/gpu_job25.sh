#!/bin/bash
# round-2 GPU pass W: what slowed the short causal shapes since r02a?  A/B of the boundary changes against the r02a kernel
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
LIBS="default $V/libhead.so $V/libnownext.so $V/libepid.so $V/libnownext_epid.so $V/libr02a.so"
for S in 1024 512 2048; do
echo "--- S=$S causal" >> $O/g25_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 400 python scripts/tune_prefill.py $LIBS >> $O/g25_tune.txt 2>&1
done
echo "--- S=1024" >> $O/g25_tune.txt
TUNE_S=1024 timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g25_tune.txt 2>&1
grep -v Traceback $O/g25_tune.txt | cut -c1-250
