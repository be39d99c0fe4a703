#!/bin/bash
# one GPU: parity tests, bench (Xvo through TMA vs generic), timeline
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_e.log
for G in 0 1; do
MPCC_XVO_GENERIC=$G timeout 600 python bench.py --steps 20 --warmup 5 2> gpurun_out/bench_n1_xvog$G.err | grep '^{' > gpurun_out/bench_n1_xvog$G.json; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n1_xvog$G.json"))
print("xvo_generic=$G", round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['ms_per_step'],3), 'parity', d['parity_vs_fixture'], {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
PY
done
timeout 600 python tools/timeline.py > gpurun_out/timeline_n1.log 2>&1; echo "timeline rc=$?"; grep -v -i warn gpurun_out/timeline_n1.log | tail -5
