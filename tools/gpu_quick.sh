#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo rc=$?
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_q.json")); r=d["roofline"]
print(d["ms_per_step"], d["value"], r["achieved"], r["yvo_gemm"]["achieved"], r["stages_ms_per_step"])
PY
python tools/sweep.py --out gpurun_out/sweep.json > gpurun_out/sweep.log 2>&1; echo "sweep rc=$?"; cut -c1-330 gpurun_out/sweep.log
