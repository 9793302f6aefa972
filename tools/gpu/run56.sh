#!/bin/bash
# generator_13_3 / 13_4 zoo entries on the GPU (structure, D(x), a finite step) + the GELU / pixnorm kernel tests again
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_other_configs.py tests/test_gpu_kernels.py -q -m gpu -p no:cacheprovider --durations=6 \
  -k "(zoo_cuda and (tree13 or tree10_2)) or pixnorm or pointwise" 2>&1 | tail -25 > gpurun_out/r56_tree13.txt
tail -14 gpurun_out/r56_tree13.txt
