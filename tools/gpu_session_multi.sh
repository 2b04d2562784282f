#!/bin/bash
# Multi-GPU session (gpurun --gpus N): the bench line at N ranks (c4 visibility range-sharded, c5 point-sharded,
# partition with slab-sharded A1).  usage: bash tools/gpu_session_multi.sh <tag> <N> [extra bench args]
tag=$1; n=$2; shift 2
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $out/${tag}_smi.txt 2>&1
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $n --steps 5 --warmup 3 "$@" > $out/${tag}_bench_n$n.json 2> $out/${tag}_bench_n$n.err
echo "rc=$?" >> $out/${tag}_bench_n$n.err
tail -c 1500 $out/${tag}_bench_n$n.err
head -c 1500 $out/${tag}_bench_n$n.json
