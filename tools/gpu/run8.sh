cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests/test_graphed_step.py -q -m gpu 2>&1 | grep -E "passed|failed|^E |eager vs" | head -20
python tools/one_conv.py 96 96 5 64 1 fprop && ncu --set full --clock-control none --import-source on -k regex:tc_conv_halo -s 1 -c 2 -f -o gpurun_out/r02_conv_fprop python tools/one_conv.py 96 96 5 64 1 fprop > gpurun_out/ncu_fprop.log 2>&1; echo "ncu fprop rc=$?"
python tools/one_conv.py 96 96 5 64 1 wgrad && ncu --set full --clock-control none --import-source on -k regex:tc_wgrad_ring -s 1 -c 2 -f -o gpurun_out/r02_conv_wgrad python tools/one_conv.py 96 96 5 64 1 wgrad > gpurun_out/ncu_wgrad.log 2>&1; echo "ncu wgrad rc=$?"
python tools/one_conv.py 64 64 3 64 1 fprop && ncu --set full --clock-control none --import-source on -k regex:tc_conv_kernel -s 1 -c 2 -f -o gpurun_out/r02_conv_d64 python tools/one_conv.py 64 64 3 64 1 fprop > gpurun_out/ncu_d64.log 2>&1; echo "ncu d64 rc=$?"
GANB200_BENCH_BATCH=256 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 30000 -c 3000 --csv --log-file gpurun_out/r02_launches_b256_eager.csv python bench.py --graphs 0 --steps 1 --warmup 1 --no-cpu --second-loop '' > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
SANITIZE_TIMEOUT=600 bash tools/sanitize.sh memcheck "grouped or (tcgen05 and fprop)" 2>&1 | tail -8
