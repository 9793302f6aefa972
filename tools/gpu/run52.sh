#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "bmm or softmax_rows or self_attention" 2>&1 | tail -4
timeout 900 python -m pytest tests/test_other_configs.py -q -m gpu -x -k "attn9_3" 2>&1 | tail -6
timeout 1800 python -m pytest tests -q -m gpu > gpurun_out/t_gpu_all5.log 2>&1; echo "full rc=$?"; tail -4 gpurun_out/t_gpu_all5.log
