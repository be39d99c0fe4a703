#!/bin/bash
# N-GPU pass (N = $1): multi-rank parity (fixtures, full shape), bench with the parity flag, CC ladder sharding check
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/mgpu_check.py --full > gpurun_out/mgpu_check_n$N.log 2>&1; echo "mgpu_check rc=$?"
grep -cE " OK" gpurun_out/mgpu_check_n$N.log; grep -E "MISMATCH|Error|error|full shape" gpurun_out/mgpu_check_n$N.log | tail -8
for MODE in ${MODES:-auto}; do
MPCC_JK_EXCHANGE=$MODE timeout 400 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/bench_n${N}_$MODE.err | grep '^{' > gpurun_out/bench_n${N}_$MODE.json; echo "bench $MODE rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_n${N}_$MODE.json"))
    print("$MODE", round(d['ms_per_step'],3), 'ms', round(d['value'],1), 'TF parity', d['parity_vs_fixture'], d['parity_detail']['max_abs_dE_diff'], d['config']['jk_exchange'], {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/bench_n${N}_$MODE.err").read()[-1500:])
PY
done
timeout 400 $TR --master-port 29513 tools/cc_mgpu_check.py > gpurun_out/cc_mgpu_n$N.log 2>&1; echo "cc_mgpu rc=$?"; grep -E "MISMATCH|^\{|rror" gpurun_out/cc_mgpu_n$N.log | tail -5
