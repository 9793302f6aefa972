#!/bin/bash
# generator_13 (tree13) on the GPU, fp32 engine
mkdir -p gpurun_out
timeout 80 python -m pytest tests/test_other_configs.py -q -m gpu -p no:cacheprovider --durations=3 -k "zoo_cuda and tree13 and not tree13_" 2>&1 | tail -25 > gpurun_out/r58_tree13.txt
tail -12 gpurun_out/r58_tree13.txt
