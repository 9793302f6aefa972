set -x
cd $GRAFT_REPO_ROOT
timeout 600 python tools/gpu/diag_graph.py 2>&1 | tail -4
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "grouped" 2>&1 | tail -15
timeout 900 python -m pytest tests/test_golden_final_pair.py tests/test_style_layer.py -q -m gpu 2>&1 | tail -8
timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_grouped.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_grouped.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d['peak_mem_gib'])"
timeout 600 python tools/kernel_profile.py 256 ganR2 > gpurun_out/cupti_grouped.txt 2>&1; head -40 gpurun_out/cupti_grouped.txt | cut -c1-140
