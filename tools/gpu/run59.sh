#!/bin/bash
# generator_12 (tree12) on the GPU, fp32 engine, + the philox noise path of generator_12 / 13
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_other_configs.py -q -m gpu -p no:cacheprovider --durations=3 -k "zoo_cuda and tree12" 2>&1 | tail -25 > gpurun_out/r59_tree12.txt
tail -8 gpurun_out/r59_tree12.txt
timeout 40 python tools/gpu/philox_zoo.py 2>&1 | tail -4 | tee -a gpurun_out/r59_tree12.txt
