#!/bin/bash
# generator_13_1 (tree13_1) on the GPU, fp32 engine
mkdir -p gpurun_out
timeout 30 python -m pytest tests/test_other_configs.py -q -m gpu -p no:cacheprovider -k "zoo_cuda and tree13_1" 2>&1 | tail -12 > gpurun_out/r60_tree13_1.txt
tail -6 gpurun_out/r60_tree13_1.txt
