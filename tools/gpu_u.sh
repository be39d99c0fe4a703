#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
timeout 600 python tools/u_bench.py > gpurun_out/u_bench.log 2>&1; echo "u_bench rc=$?"; tail -5 gpurun_out/u_bench.log | cut -c1-400
