#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
for d in 0 1 2 3 4 7; do echo "== GANB200_FLAT_DEBUG=$d"; GANB200_FLAT_DEBUG=$d timeout 300 python tools/flat_bench.py 2>&1 | sed -n 3,7p; done
