#!/bin/bash
# round-2 GPU pass S: step-level causal traces
O=gpurun_out; mkdir -p $O
V=mini_flash_attention_b200/variants
for S in 1024 4096; do
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py $S causal > $O/g20_trace_s${S}c.txt 2>&1; tail -60 $O/g20_trace_s${S}c.txt | cut -c1-330
done
MFA_B200_SHIM=py MFA_B200_LIB=$V/libtrace.so timeout 200 python scripts/trace_prefill.py 4096 > $O/g20_trace_s4096.txt 2>&1; tail -16 $O/g20_trace_s4096.txt | cut -c1-330
