cd $GRAFT_REPO_ROOT
timeout 600 python tools/scatter_bench.py 256 2>&1 | tail -16
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "scatter or conv_family or fused_style" 2>&1 | tail -4
timeout 600 python -m pytest tests/test_golden_final_pair.py -q -m gpu 2>&1 | tail -30 | cut -c1-300
timeout 900 python bench.py --steps 6 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_scatter1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_scatter1.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d['peak_mem_gib'], d['roofline']['frac'])"
