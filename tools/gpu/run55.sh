#!/bin/bash
# new kernels (pixnorm, GELU, depthwise 5x5/7x7, 9x9 / 5x5-stride-2 convs) and the tree-generator zoo on the GPU
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_kernels.py -q -m gpu -p no:cacheprovider \
  -k "pixnorm or pointwise or depthwise or (conv_family_fp32 and (case33 or case34))" 2>&1 | tail -15 > gpurun_out/r55_kernels.txt
tail -3 gpurun_out/r55_kernels.txt
timeout 170 python -m pytest tests/test_other_configs.py -q -m gpu -p no:cacheprovider --durations=15 \
  -k "zoo_cuda and tree" 2>&1 | tail -60 > gpurun_out/r55_zoo.txt
tail -25 gpurun_out/r55_zoo.txt
