#!/bin/bash
# one GPU: parity tests, bench with the folded-in / stored extrapolated vector, kernel timeline
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_d.log
for IMP in auto 0; do
MPCC_DIIS_IMPLICIT=$IMP timeout 600 python bench.py --steps 20 --warmup 5 2> gpurun_out/bench_n1_imp$IMP.err | grep '^{' > gpurun_out/bench_n1_imp$IMP.json; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n1_imp$IMP.json"))
print("$IMP", round(d['ms_per_step'],3), 'ms parity', d['parity_vs_fixture'], {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})
PY
done
timeout 600 python tools/timeline.py > gpurun_out/timeline_n1.log 2>&1; echo "timeline rc=$?"; tail -25 gpurun_out/timeline_n1.log
