#!/bin/bash
# Scaling pass on an 8-GPU box: parity at 8 ranks, then the bench at 2, 4, 8 ranks.
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
    tools/mgpu_check.py --full > gpurun_out/mgpu_check_8.log 2>&1; echo "mgpu_check rc=$?"
grep -E "OK|MISMATCH|Error|error" gpurun_out/mgpu_check_8.log | tail -12
for N in 2 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
    bench.py --gpus $N --steps 5 --warmup 3 2> gpurun_out/bench_n$N.err | grep '^{' > gpurun_out/bench_n$N.json; echo "bench $N rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_n$N.json"))
    print($N, round(d['ms_per_step'],2), round(d['value'],1), round(d['e2e']['ms_per_step'],2), {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/bench_n$N.err").read()[-1500:])
PY
done
