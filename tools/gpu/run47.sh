#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py -q -m gpu -x 2>&1 | tail -3
for f in 0 1; do echo "== GANB200_FOLD_WARP_PLANES=$f"; GANB200_FOLD_WARP_PLANES=$f timeout 300 python tools/modbwd_bench.py 2>&1 | head -6; done
