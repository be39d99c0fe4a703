#!/bin/bash
# N-GPU final pass (N = $1): bench, config-4 sweep and parity, CC ladder sharding, fixtures, timeline
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/bench_final_n$N.err | grep '^{' > gpurun_out/bench_final_n$N.json; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_final_n$N.json"))
    print(round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['ms_per_step'],3), round(d['value'],1), 'TF parity', d['parity_vs_fixture'], d['parity_detail']['max_abs_dE_diff'], d['config']['jk_exchange'], d['config'].get('xoo'), {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
except Exception as e:
    print("ERR", e); print(open("gpurun_out/bench_final_n$N.err").read()[-1500:])
PY
timeout 500 $TR --master-port 29515 tools/sweep.py --out gpurun_out/sweep_config4_n$N.json --steps 5 --warmup 2 --shapes 80,848,2224 > gpurun_out/sweep_config4_n$N.log 2>&1; echo "sweep rc=$?"; grep '^{' gpurun_out/sweep_config4_n$N.log | cut -c1-900
timeout 600 $TR --master-port 29516 tools/mgpu_check.py --shape=80,848,2224 --only-shape > gpurun_out/mgpu_check_config4_n$N.log 2>&1; echo "config4 parity rc=$?"; grep -E "full shape|rror" gpurun_out/mgpu_check_config4_n$N.log | tail -3
timeout 400 $TR --master-port 29513 tools/cc_mgpu_check.py > gpurun_out/cc_mgpu_n$N.log 2>&1; echo "cc_mgpu rc=$?"; grep -E "MISMATCH|^\{|rror" gpurun_out/cc_mgpu_n$N.log | tail -3
timeout 300 $TR --master-port 29514 tools/timeline.py > gpurun_out/timeline_n$N.log 2>&1; echo "timeline rc=$?"; grep "^{\|^gap" gpurun_out/timeline_n$N.log | tail -4
timeout 600 $TR --master-port 29511 tools/mgpu_check.py --full > gpurun_out/mgpu_check_n$N.log 2>&1; echo "mgpu_check rc=$?"
grep -cE " OK" gpurun_out/mgpu_check_n$N.log; grep -E "MISMATCH|Error|error|full shape" gpurun_out/mgpu_check_n$N.log | tail -4
