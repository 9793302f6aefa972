set -x
cd $GRAFT_REPO_ROOT
timeout 1800 python -m pytest tests -q -m gpu > gpurun_out/t_gpu_all2.log 2>&1; echo "gpu tests rc=$?"
tail -15 gpurun_out/t_gpu_all2.log
timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_merged.log 2>&1; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_merged.log
