#!/bin/bash
# round-2 GPU pass G (2 GPUs): torchrun parity tests + bench at N=2 (graph-replayed KV-sharded decode, head-sharded prefill)
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/g8_topo.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -q -s > $O/g8_pytest_multigpu.log 2>&1; tail -4 $O/g8_pytest_multigpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --no-cpu-baseline > $O/g8_bench_n2.json 2> $O/g8_bench_n2.err; cat $O/g8_bench_n2.json; tail -5 $O/g8_bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --impl reference > $O/g8_bench_ref_n2.json 2> $O/g8_bench_ref_n2.err; cat $O/g8_bench_ref_n2.json | cut -c1-400
