#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench, ncu launch list, ncu full capture of the two DMMA GEMMs.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
python tools/profile_step.py 64 512 1536 2 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python tools/profile_step.py 64 512 1536 2 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -c 2 -o gpurun_out/prof_gemm \
    python tools/profile_step.py 64 512 1536 1 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_diis_push|k_diis_combine|k_dress_finalize' -c 4 -o gpurun_out/prof_mem \
    python tools/profile_step.py 64 512 1536 3 > gpurun_out/ncu_full_mem.log 2>&1
echo "ncu mem rc=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log | tail -2; cat gpurun_out/bench.json | cut -c1-600
