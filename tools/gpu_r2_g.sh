#!/bin/bash
# one GPU: parity tests with the flat-K Yvo form, sweep of awkward occupied counts (flat vs 16-wide steps)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_g.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_g.log | cut -c1-300
for FL in 0 -1; do
MPCC_YVO_FLAT=$FL timeout 900 python tools/sweep.py --out gpurun_out/sweep_flat$FL.json --steps 10 --warmup 3 --shapes 21,243,654 22,243,654 33,256,768 42,512,1536 > gpurun_out/sweep_flat$FL.log 2>&1; echo "sweep flat=$FL rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/sweep_flat$FL.json"))
    rows = d if isinstance(d, list) else d.get('results', [])
    for r in rows:
        print("flat=$FL", (r.get('nocc'), r.get('nvir'), r.get('naux')), round(r.get('ms_per_updated_amps',0),3), 'ms', 'exec TF', round(r.get('tflops_executed',0),2), 'yvo TF', round(r.get('yvo_tflops',0),2), 'frac nominal', round(r.get('frac_of_nominal_executed',0),3))
except Exception as e:
    print("ERR", e); print(open("gpurun_out/sweep_flat$FL.log").read()[-1200:])
PY
done
