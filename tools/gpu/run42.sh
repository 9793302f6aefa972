#!/bin/bash
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py -q -m gpu -x 2>&1 | tail -4
for c in 0 1; do echo "== GANB200_WGRAD_M64=$c"; GANB200_WGRAD_M64=$c timeout 600 python tools/conv_bench.py 2>&1 | grep "wgrad" | cut -c1-100; done
