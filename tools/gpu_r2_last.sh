#!/bin/bash
# last check of the tree as committed: GPU tests, smoke(), one-GPU bench (both arms)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_last.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu_last.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_last.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke_last.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/bench_last_default.json 2> gpurun_out/bench_last_default.err; echo "bench default rc=$?"
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/bench_last_default.json") if l.startswith("{")][-1])
print(d["steps"], d["warmup"], round(d["ms_per_step"], 3), "ms", round(d["value"], 3), d["unit"], "e2e", round(d["e2e"]["value"], 3), "parity", d["parity_vs_fixture"], "launches", d["gpu_launches"], "frac", round(d["roofline"]["frac"], 4))
PY
