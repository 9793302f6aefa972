cd $GRAFT_REPO_ROOT
echo v1; GANB200_PACK_V2=0 python tools/pack_bench.py 2>&1 | tail -7
echo v2; GANB200_PACK_V2=1 python tools/pack_bench.py 2>&1 | tail -7
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "conv_family or scatter or fused_style or dominant" 2>&1 | tail -3
