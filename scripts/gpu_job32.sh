#!/bin/bash
# ncu captures of the prefill kernel as it stands at the end of round 2 (tag r02d): headline, causal, S=512
O=gpurun_out; mkdir -p $O/profiles
TAG=r02d
cp profiles/ncu_summary.json $O/profiles/ncu_summary.json
for CASE in prefill prefill_causal prefill_s512; do
  python scripts/profile_cases.py $CASE 2 > $O/plain_case.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:prefill_fwd -s 2 -c 1 -f -o $O/${CASE}_$TAG python scripts/profile_cases.py $CASE 2 > $O/ncu_$CASE.log 2>&1
  echo "ncu $CASE rc=$?"
done
python scripts/ncu_summarize.py $TAG --out $O/profiles prefill prefill_causal prefill_s512 > $O/g32_summarize.log 2>&1; tail -5 $O/g32_summarize.log
rm -f $O/*.ncu-rep
ls $O/profiles
