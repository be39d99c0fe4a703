#!/bin/bash
# Multi-GPU pass: N-rank parity vs golden + sharded-vs-single at the target shape, then the bench at N ranks.
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    tools/mgpu_check.py --full > gpurun_out/mgpu_check_$N.log 2>&1; echo "mgpu_check rc=$?"
tail -12 gpurun_out/mgpu_check_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$N.err; cut -c1-400 gpurun_out/bench_n$N.json
