cd $GRAFT_REPO_ROOT
nvidia-smi -L
timeout 900 python -m pytest tests/test_dist_nccl.py -q -m gpu -x 2>&1 | tail -15
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_2gpu.log 2>&1; echo "bench2 rc=$?"; tail -3 gpurun_out/bench_2gpu.log | cut -c1-1500
