#!/bin/bash
# one GPU, evidence pass: parity tests, driver-configuration bench (both arms), ncu launch list and full captures
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu_final.log | cut -c1-200
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_reference.json 2> gpurun_out/bench_final_reference.err; echo "reference arm rc=$?"
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/launches_final.csv \
    $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
# per step the TMA kernel is launched as Xvo <2,2>, XT|JT <0,4>, eris/t2 <1,2>, Yvo <0,4>: skip two whole steps
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -s 8 -c 4 -o gpurun_out/prof_final_gemms \
    $CMD > gpurun_out/ncu_gemms.log 2>&1; echo "ncu gemms rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_diis_push|k_dress_finalize|k_t1_update|k_gemv_rows" -s 12 -c 4 -o gpurun_out/prof_final_elem \
    $CMD > gpurun_out/ncu_elem.log 2>&1; echo "ncu elem rc=$?"
for f in prof_final_gemms prof_final_elem; do ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null; done
rm -f gpurun_out/prof_final_elem.ncu-rep
timeout 600 python tools/cc_bench.py > gpurun_out/cc_bench_final.log 2>&1; echo "cc_bench rc=$?"; tail -3 gpurun_out/cc_bench_final.log | cut -c1-400
python - <<PY
import json
for f in ("gpurun_out/bench_final_n1.json", "gpurun_out/bench_final_reference.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["ms_per_step"], 3), "ms", round(d["value"], 4), d["unit"], "e2e", d.get("e2e", {}).get("value"), "parity", d.get("parity_vs_fixture"),
              "frac", d.get("roofline", {}).get("frac"), "yard", d.get("roofline", {}).get("peak"))
    except Exception as e:
        print(f, "ERR", e)
PY
