#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline > gpurun_out/bench_sym.json 2> gpurun_out/bench_sym.err; echo rc=$?
MPCC_T2_SYM=0 python bench.py --no-cpu-baseline > gpurun_out/bench_nosym.json 2> gpurun_out/bench_nosym.err; echo rc=$?
python - <<'PY'
import json
for f in ("sym","nosym"):
    try:
        d=json.load(open(f"gpurun_out/bench_{f}.json")); r=d["roofline"]
        print(f, d["ms_per_step"], d["value"], r["achieved"], r.get("algorithmic_tflops"), r["yvo_gemm"]["achieved"], r["stages_ms_per_step"])
    except Exception as e: print(f, "ERR", e, open(f"gpurun_out/bench_{f}.err").read()[-1200:])
PY
