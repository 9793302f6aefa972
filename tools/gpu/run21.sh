cd $GRAFT_REPO_ROOT
nvidia-smi -L | wc -l
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_8gpu.log 2>&1; echo "bench8 rc=$?"; tail -1 gpurun_out/bench_8gpu.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('nccl_calls_per_step'), d.get('bce_in_range'), d['peak_mem_gib'])" || tail -20 gpurun_out/bench_8gpu.log | cut -c1-300
