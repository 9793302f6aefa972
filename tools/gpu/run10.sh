cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_final_pair_blockwise.py -q -m gpu -s 2>&1 | grep -E "passed|failed|worst|^E " | cut -c1-400 | head -20
GANB200_PACK_CACHE=0 python tools/one_conv.py 96 96 5 64 1 fprop && GANB200_PACK_CACHE=0 ncu --set full --clock-control none --import-source on -k regex:pack_act_vec -s 1 -c 2 -f -o gpurun_out/r02_pack python tools/one_conv.py 96 96 5 64 1 fprop > gpurun_out/ncu_pack.log 2>&1; echo "ncu pack rc=$?"
GANB200_PACK_CACHE=0 ncu --set full --clock-control none --import-source on -k regex:fold_scale -s 0 -c 1 -f -o gpurun_out/r02_fold python tools/one_modbwd.py > gpurun_out/ncu_fold.log 2>&1; echo "ncu fold rc=$?"
