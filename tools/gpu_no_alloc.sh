#!/bin/bash
# mpccb_dot_static: its test alone in a fresh process, then the whole GPU suite (the new file is collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/pytest_no_alloc_alone.log | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_with_no_alloc.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/pytest_gpu_with_no_alloc.log | cut -c1-300
