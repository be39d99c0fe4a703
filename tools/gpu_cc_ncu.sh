#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_rmpccsd.py -m gpu -x -q > gpurun_out/pytest_cc.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_cc.log
python tools/cc_bench.py 32,128,384 > gpurun_out/cc_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_cc.csv \
    python tools/cc_bench.py 32,128,384 > gpurun_out/ncu_cc.log 2>&1
echo "ncu rc=$?"
