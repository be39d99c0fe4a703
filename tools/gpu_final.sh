#!/bin/bash
# sweep refresh + compute-sanitizer memcheck on small cases
mkdir -p gpurun_out
python tools/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1 &&
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/sanitize_small.py > gpurun_out/sanitize_memcheck.log 2>&1
echo "memcheck rc=$?"; tail -6 gpurun_out/sanitize_memcheck.log
python tools/sweep.py --out gpurun_out/sweep.json > gpurun_out/sweep.log 2>&1; echo "sweep rc=$?"
python - <<'PY'
import json
for r in json.load(open("gpurun_out/sweep.json")):
    if "tflops" in r: print(r["nocc"], r["nvir"], r["naux"], round(r["ms_per_updated_amps"],2), round(r["tflops"],1), round(r["eris_t2_tflops"],1), round(r["yvo_tflops"],1))
    else: print(r)
PY
