cd $GRAFT_REPO_ROOT
GANB200_MODCONV_RED=1 timeout 600 python tools/scatter_bench.py 256 2>&1 | tail -16
GANB200_MODCONV_RED=1 timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "scatter or fused_style" 2>&1 | tail -4
