#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "tcgen05 or conv" 2>&1 | tail -3
for cfg in "0" "2"; do echo "== mt=$cfg"; GANB200_FLAT_MT=$cfg timeout 300 python tools/flat_bench.py 2>&1 | sed -n 3,14p; done
