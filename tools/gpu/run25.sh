#!/bin/bash
# 2 GPUs: peer exchange + fused sync BN: tests, N=2 bench
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dist_nccl.py -q -m gpu -s > gpurun_out/t_peer.log 2>&1
echo "peer rc=$?"; tail -15 gpurun_out/t_peer.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 6 --warmup 3 --no-cpu > gpurun_out/bench_2gpu_peer.log 2>gpurun_out/bench_2gpu_peer.err
echo "bench2 rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/bench_2gpu_peer.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('nccl_calls_per_step'), d.get('peer_calls_per_step'), d.get('peer_fused_per_step'), d.get('bce_in_range'), d.get('losses'))
PY
tail -5 gpurun_out/bench_2gpu_peer.err
