#!/bin/bash
# parity tests, driver-configuration bench, launch list + full captures of the two GEMMs and the DIIS push (raw CSV)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench20.json 2> gpurun_out/bench20.err; echo "bench20 rc=$?"
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/launches_bench.csv \
    $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -s 20 -c 2 -o gpurun_out/prof_gemms \
    $CMD > gpurun_out/ncu_gemms.log 2>&1; echo "ncu gemms rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_diis_push -s 9 -c 1 -o gpurun_out/prof_diis \
    $CMD > gpurun_out/ncu_diis.log 2>&1; echo "ncu diis rc=$?"
for f in prof_gemms prof_diis; do ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null; done
cut -c1-300 gpurun_out/bench20.json
