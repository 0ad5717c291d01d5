#!/bin/bash
# round-2 GPU pass H: full parity suite, bf16 error sweep, bench (both arms), ncu launch list + per-kernel captures
O=gpurun_out; mkdir -p $O $O/profiles
TAG=r02a
timeout 1500 python -m pytest tests -m gpu -q -s > $O/g9_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g9_pytest.log; grep -E "passed|failed|reference decode rows|^FAILED" $O/g9_pytest.log | tail -8
timeout 600 python scripts/bf16_error_vs_s.py > $O/profiles/bf16_error_vs_s_$TAG.jsonl 2> $O/g9_bf16.err; tail -3 $O/profiles/bf16_error_vs_s_$TAG.jsonl
timeout 600 python bench.py --impl reference > $O/profiles/bench_ref_$TAG.json 2> $O/g9_bench_ref.err
timeout 900 python bench.py > $O/profiles/bench_ours_$TAG.json 2> $O/g9_bench.err; cat $O/profiles/bench_ours_$TAG.json
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-sustained --no-libs"
$SHORT > $O/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$TAG.csv $SHORT > $O/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
for CASE in prefill prefill_causal prefill_s512 varlen decode decode_kv512 decode_gqa_kv1024 decode_fp8 combine xchg append; do
  case $CASE in
    prefill*|varlen) K=prefill_fwd;; decode*) K=decode_splitkv;; combine) K=combine_kernel;; xchg) K=xchg_merge;; append) K=kvcache_append;;
  esac
  python scripts/profile_cases.py $CASE 2 > $O/plain_case.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$K -s 2 -c 1 -f -o $O/${CASE}_$TAG python scripts/profile_cases.py $CASE 2 > $O/ncu_$CASE.log 2>&1
  echo "ncu $CASE rc=$?"
done
python scripts/ncu_summarize.py $TAG --out $O/profiles prefill prefill_causal prefill_s512 varlen decode decode_kv512 decode_gqa_kv1024 decode_fp8 combine xchg append > $O/g9_summarize.log 2>&1; tail -15 $O/g9_summarize.log
rm -f $O/*.ncu-rep
ls -la $O/profiles
