#!/bin/bash
# N-rank parity + bench for each way of exchanging the dressed integrals (fused multicast / p2p stores / NCCL)
N=${1:-2}
mkdir -p gpurun_out
for MODE in auto p2p nccl; do
export MPCC_JK_EXCHANGE=$MODE
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 \
    tools/mgpu_check.py --full > gpurun_out/mgpu_${MODE}_$N.log 2>&1; echo "$MODE mgpu_check rc=$?"
grep -E "OK|MISMATCH|Error|error" gpurun_out/mgpu_${MODE}_$N.log | tail -9
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus $N --steps 5 --warmup 3 2> gpurun_out/bench_${MODE}_n$N.err | grep '^{' > gpurun_out/bench_${MODE}_n$N.json; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_${MODE}_n$N.json"))
    print("$MODE", d['config']['jk_exchange'], round(d['ms_per_step'],2), round(d['value'],1), round(d['e2e']['ms_per_step'],2), {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/bench_${MODE}_n$N.err").read()[-1500:])
PY
done
