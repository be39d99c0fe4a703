#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 600 python tools/cc_bench.py > gpurun_out/cc_bench.log 2>&1; echo "cc_bench rc=$?"; tail -8 gpurun_out/cc_bench.log | cut -c1-400
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench20.json 2> gpurun_out/bench20.err; echo "bench20 rc=$?"
cut -c1-300 gpurun_out/bench20.json
