#!/bin/bash
# Evidence pass: parity tests, bench, then ncu on the SAME bench command (launch list + full capture of the DMMA GEMMs).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_bench.csv \
    $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
# eris/t2 kernel: 9th match of the regex (5th iteration's eris); Yvo: 10th match, captured in its own ncu run
ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -s 8 -c 1 -o gpurun_out/prof_eris \
    $CMD > gpurun_out/ncu_eris.log 2>&1; echo "ncu eris rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -s 9 -c 1 -o gpurun_out/prof_yvo \
    $CMD > gpurun_out/ncu_yvo.log 2>&1; echo "ncu yvo rc=$?"
cut -c1-400 gpurun_out/bench_full.json
