#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python tools/sweep.py --out gpurun_out/sweep_odd.json --steps 10 --warmup 3 --shapes 21,243,654 22,243,654 33,256,768 32,256,768 > gpurun_out/sweep_odd.log 2>&1; echo "sweep rc=$?"; cat gpurun_out/sweep_odd.log | cut -c1-700
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench20.json 2> gpurun_out/bench20.err; echo "bench20 rc=$?"
cut -c1-300 gpurun_out/bench20.json
