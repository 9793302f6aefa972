#!/bin/bash
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py -q -m gpu -x > gpurun_out/t_flat.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t_flat.log
timeout 600 python tools/flat_bench.py > gpurun_out/flat_bench.log 2>&1; echo "bench rc=$?"; cat gpurun_out/flat_bench.log
timeout 600 python tools/modbwd_bench.py 2>&1 | head -8
