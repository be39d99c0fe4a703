#!/bin/bash
# Round-2 first GPU pass: parity tests (new BASELINE-shape pins), the driver's bench configuration, a 40-step one-rank
# trace for the bench-side parity fixture, and the reference arm on the staged reference files.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 40 --warmup 5 --no-cpu-baseline --write-trace gpurun_out/bench_trace_64_512_1536.json > gpurun_out/bench40.json 2> gpurun_out/bench40.err; echo "bench40 rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/bench20.json 2> gpurun_out/bench20.err; echo "bench20 rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-600 gpurun_out/bench20.json; tail -3 gpurun_out/bench20.err
