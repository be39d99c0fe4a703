#!/bin/bash
# Evidence pass: parity tests, bench, then ncu on the SAME bench command (launch list + full capture of the DMMA GEMMs).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_bench.csv \
    $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -s 6 -c 2 -o gpurun_out/prof_gemm2 \
    $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -4 gpurun_out/ncu_full.log
cut -c1-300 gpurun_out/bench_plain.json
