#!/bin/bash
# round-2 GPU pass V (final state of round 2): full parity suite, bf16 error sweep, bench (both arms), ncu launch list + per-kernel captures
O=gpurun_out; mkdir -p $O $O/profiles
TAG=r02b
timeout 1500 python -m pytest tests -m gpu -q -s > $O/g24_pytest.log 2>&1; echo "pytest rc=$?" >> $O/g24_pytest.log; grep -E "passed|failed|reference decode rows|^FAILED" $O/g24_pytest.log | tail -8
timeout 600 python scripts/bf16_error_vs_s.py > $O/profiles/bf16_error_vs_s_$TAG.jsonl 2> $O/g24_bf16.err; tail -3 $O/profiles/bf16_error_vs_s_$TAG.jsonl
timeout 600 python bench.py --impl reference > $O/profiles/bench_ref_$TAG.json 2> $O/g24_bench_ref.err
timeout 900 python bench.py > $O/profiles/bench_ours_$TAG.json 2> $O/g24_bench.err; cat $O/profiles/bench_ours_$TAG.json
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
python scripts/ncu_summarize.py $TAG --out $O/profiles prefill prefill_causal prefill_s512 varlen decode decode_kv512 decode_gqa_kv1024 decode_fp8 combine xchg append > $O/g24_summarize.log 2>&1; tail -15 $O/g24_summarize.log
timeout 900 python scripts/sweep.py > $O/profiles/sweep_$TAG.jsonl 2> $O/g24_sweep.err; tail -3 $O/profiles/sweep_$TAG.jsonl
timeout 600 python scripts/decode_splits_gqa.py > $O/profiles/decode_splits_$TAG.txt 2>&1
MFA_B200_SHIM=py MFA_B200_LIB=mini_flash_attention_b200/variants/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/profiles/trace_prefill_$TAG.txt 2>&1
MFA_B200_SHIM=py MFA_B200_LIB=mini_flash_attention_b200/variants/libtrace.so timeout 200 python scripts/trace_prefill.py 512 > $O/profiles/trace_prefill_s512_$TAG.txt 2>&1
MFA_B200_SHIM=py MFA_B200_LIB=mini_flash_attention_b200/variants/libtrace.so timeout 200 python scripts/trace_prefill.py 1024 causal > $O/profiles/trace_prefill_s1024_causal_$TAG.txt 2>&1
rm -f $O/*.ncu-rep
ls -la $O/profiles
