#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -q -m gpu > gpurun_out/t_gpu_all6.log 2>&1; echo "full rc=$?"; tail -4 gpurun_out/t_gpu_all6.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
