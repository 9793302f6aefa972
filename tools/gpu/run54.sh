#!/bin/bash
# depthwise 3/5/7 kernels + the tree critic (tree6) on the GPU
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_kernels.py tests/test_other_configs.py tests/test_zz_no_alloc.py -q -m gpu -k "depthwise or tree6 or alloc" 2>&1 | tail -15 > gpurun_out/r54_tests.txt
timeout 90 python tools/gpu/tree6_stats.py fp32 > gpurun_out/r54_tree6_fp32.txt 2>&1
timeout 60 python tools/gpu/tree6_stats.py auto > gpurun_out/r54_tree6_auto.txt 2>&1
tail -5 gpurun_out/r54_tests.txt
