#!/bin/bash
# N-GPU pass with the kernel timeline (N = $1); VARIANTS = space-separated "NAME:ENV=VAL,ENV=VAL" bench variants
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/mgpu_check.py --full > gpurun_out/mgpu_check_n$N.log 2>&1; echo "mgpu_check rc=$?"
grep -cE " OK" gpurun_out/mgpu_check_n$N.log; grep -E "MISMATCH|Error|error|full shape" gpurun_out/mgpu_check_n$N.log | tail -8
for VAR in ${VARIANTS:-auto:}; do
NAME=${VAR%%:*}; ENVS=${VAR#*:}
env $(echo $ENVS | tr ',' ' ') timeout 400 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/bench_n${N}_$NAME.err | grep '^{' > gpurun_out/bench_n${N}_$NAME.json; echo "bench $NAME rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_n${N}_$NAME.json"))
    print("$NAME", round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['ms_per_step'],3), round(d['value'],1), 'TF parity', d['parity_vs_fixture'], d['parity_detail']['max_abs_dE_diff'], d['config']['jk_exchange'], d['config'].get('xoo'), {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/bench_n${N}_$NAME.err").read()[-1500:])
PY
done
timeout 400 $TR --master-port 29514 tools/timeline.py > gpurun_out/timeline_n$N.log 2>&1; echo "timeline rc=$?"; grep -v -i "warn\|\*\*\*\|OMP" gpurun_out/timeline_n$N.log | tail -8
