#!/bin/bash
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "tcgen05 or conv" > gpurun_out/t_flat.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/t_flat.log
timeout 600 python tools/flat_bench.py > gpurun_out/flat_bench.log 2>&1; echo "bench rc=$?"; cat gpurun_out/flat_bench.log
