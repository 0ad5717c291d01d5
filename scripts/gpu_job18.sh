#!/bin/bash
# round-2 GPU pass Q: which of the item-boundary changes costs the causal shapes? (A/B by compile-time switches)
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
LIBS="default $V/libnownext.so $V/libnownext_qtog.so $V/libnownext_qtog_epid.so $V/libqtogether.so $V/libepidirect.so"
for S in 1024 512 4096; do
echo "--- S=$S causal" >> $O/g18_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 400 python scripts/tune_prefill.py $LIBS >> $O/g18_tune.txt 2>&1
done
echo "--- S=512" >> $O/g18_tune.txt
TUNE_S=512 timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g18_tune.txt 2>&1
grep -v Traceback $O/g18_tune.txt | cut -c1-250
