cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "scatter or conv_family or fused_style" 2>&1 | tail -12
timeout 900 python -m pytest tests/test_golden_final_pair.py tests/test_other_configs.py tests/test_final_pair_blockwise.py -q -m gpu -x 2>&1 | tail -5
for m in 0 1; do GANB200_DGRAD_SCATTER=$m timeout 900 python bench.py --steps 6 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_scatter$m.log 2>&1; echo "bench scatter=$m rc=$?"; tail -1 gpurun_out/bench_scatter$m.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d['peak_mem_gib'], d['roofline']['frac'])"; done
