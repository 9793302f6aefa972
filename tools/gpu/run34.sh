#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py -q -m gpu -x 2>&1 | tail -3
for c in 0 1; do echo "== GANB200_CPAD64=$c flat=2"; GANB200_CPAD64=$c timeout 300 python tools/flat_bench.py 2>&1 | sed -n 3,15p; done
for c in 0 1; do echo "== GANB200_CPAD64=$c conv_bench"; GANB200_CPAD64=$c timeout 600 python tools/conv_bench.py 2>&1 | grep -v "^B=\|^shape" | cut -c1-100; done
