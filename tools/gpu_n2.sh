#!/bin/bash
# 2-GPU pass: parity (equal and uneven slabs), then the bench with the column-owned and the row-owned DIIS vector.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 500 $TR --master-port 29511 tools/mgpu_check.py --full > gpurun_out/mgpu_check_2.log 2>&1; echo "mgpu_check rc=$?"
grep -E "OK|MISMATCH|Error|error" gpurun_out/mgpu_check_2.log | tail -14
for S in columns rows; do
timeout 300 $TR --master-port 29512 bench.py --gpus 2 --steps 8 --warmup 6 --diis-layout $S 2> gpurun_out/bench_n2_$S.err | grep '^{' > gpurun_out/bench_n2_$S.json; echo "bench $S rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_n2_$S.json"))
    print("$S", round(d['ms_per_step'],2), round(d['value'],1), d['config'].get('slab_bounds'), {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/bench_n2_$S.err").read()[-1500:])
PY
done
