#!/bin/bash
# 8 GPUs: peer exchange test on 8 ranks, N=8 bench
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist_nccl.py -q -m gpu -k peer > gpurun_out/t_peer8.log 2>&1
echo "peer8 rc=$?"; tail -4 gpurun_out/t_peer8.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 8 --steps 6 --warmup 3 --no-cpu > gpurun_out/bench_8gpu_peer.log 2>gpurun_out/bench_8gpu_peer.err
echo "bench8 rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/bench_8gpu_peer.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('nccl_calls_per_step'), d.get('peer_calls_per_step'), d.get('peer_fused_per_step'), d.get('bce_in_range'), d.get('losses'))
PY
tail -5 gpurun_out/bench_8gpu_peer.err
