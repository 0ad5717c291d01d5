#!/bin/bash
# round-2 GPU pass X: Q waits back at the top of the item (the per-tile wait machinery cost the MMA warp 8 % on short causal
# shapes), epilogue form chosen per launch; dependent-launch A/B for the split decode
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
timeout 1200 python -m pytest tests -q -x -m gpu > $O/g26_pytest.log 2>&1; tail -3 $O/g26_pytest.log
LIBS="default $V/libhead.so"
for S in 512 1024 4096; do
echo "--- S=$S causal" >> $O/g26_tune.txt
TUNE_S=$S TUNE_CAUSAL=1 timeout 400 python scripts/tune_prefill.py $LIBS >> $O/g26_tune.txt 2>&1
done
echo "--- S=4096" >> $O/g26_tune.txt
timeout 300 python scripts/tune_prefill.py $LIBS >> $O/g26_tune.txt 2>&1
grep -v Traceback $O/g26_tune.txt | cut -c1-250
mkdir -p $O/profiles
timeout 600 python scripts/decode_splits_gqa.py small > $O/profiles/decode_pdl_r02b.txt 2>&1
MFA_DECODE_NO_PDL=1 timeout 600 python scripts/decode_splits_gqa.py small >> $O/profiles/decode_pdl_r02b.txt 2>&1
cat $O/profiles/decode_pdl_r02b.txt
timeout 900 python scripts/sweep.py > $O/profiles/sweep_r02b.jsonl 2> $O/g26_sweep.err; grep -c . $O/profiles/sweep_r02b.jsonl
