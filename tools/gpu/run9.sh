cd $GRAFT_REPO_ROOT
timeout 1800 python -m pytest tests -q -m gpu -x 2>&1 | tail -12
timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_packcache.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_packcache.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d['peak_mem_gib'], d['roofline']['frac'])"
timeout 600 python tools/kernel_profile.py 256 ganR2 > gpurun_out/cupti_packcache.txt 2>&1; head -24 gpurun_out/cupti_packcache.txt | cut -c1-140
SANITIZE_TIMEOUT=300 timeout 400 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "grouped_style_bank_cuda_equals" 2>&1 | tail -12
