#!/bin/bash
# the kernel-level GPU tests after the depthwise / pointwise / pixnorm changes (whole file) + smoke
mkdir -p gpurun_out
timeout 95 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider -x 2>&1 | tail -6 > gpurun_out/r57_kernels_all.txt
tail -4 gpurun_out/r57_kernels_all.txt
timeout 25 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
