#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_rmpccsd.py tests/test_gpu_umpcc.py -m gpu -x -q > gpurun_out/pytest_cc.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_cc.log
timeout 600 python tools/cc_bench.py 32,128,384 32,192,576 > gpurun_out/cc_bench.log 2>&1; echo "cc_bench rc=$?"
python - <<'PY'
import json
for r in json.load(open('gpurun_out/cc_bench.json')):
    print(r['nocc'], r['nvir'], r['naux'], round(r['ms_update_amps'],1), 'ms', round(r['tflops'],2), 'TF', 'einsum total', round(r['einsum_ms_total'],1), 'profiled total', round(r['profiled_ms_total'],1))
    for t in r['top_einsums']: print('   ', t)
PY
