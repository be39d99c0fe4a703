#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
python tools/sweep.py --out gpurun_out/sweep_gen.json --shapes 21,243,654 22,243,654 40,320,960 80,848,2224 96,512,1536 > gpurun_out/sweep_gen.log 2>&1; echo "sweep rc=$?"
python - <<'PY'
import json
for r in json.load(open("gpurun_out/sweep_gen.json")):
    if "tflops" in r: print(r["nocc"], r["nvir"], r["naux"], round(r["ms_per_updated_amps"],2), round(r["tflops"],1), round(r["eris_t2_tflops"],1), round(r["yvo_tflops"],1))
    else: print(r)
PY
