#!/bin/bash
# 8 GPUs: N=8 bench, N=2 bench, and the data-parallel parity tests
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 8 --steps 8 --warmup 3 --no-cpu > gpurun_out/bench_8gpu_final.log 2>gpurun_out/bench_8gpu_final.err
echo "bench8 rc=$?"
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 4 --steps 8 --warmup 3 --no-cpu > gpurun_out/bench_4gpu_final.log 2>gpurun_out/bench_4gpu_final.err
echo "bench4 rc=$?"
CUDA_VISIBLE_DEVICES=0,1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29703 bench.py --gpus 2 --steps 8 --warmup 3 --no-cpu > gpurun_out/bench_2gpu_final.log 2>gpurun_out/bench_2gpu_final.err
echo "bench2 rc=$?"
CUDA_VISIBLE_DEVICES=0 timeout 900 python bench.py --gpus 1 --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_1gpu_samebox.log 2>/dev/null
echo "bench1 rc=$?"
python - <<'PY'
import json
for f in ('bench_1gpu_samebox','bench_2gpu_final','bench_4gpu_final','bench_8gpu_final'):
    for l in open(f'gpurun_out/{f}.log'):
        if l.startswith('{'):
            d=json.loads(l); print(f, d['value'], d['ms_per_step'], d['gpu_launches'], d.get('nccl_calls_per_step'), d.get('peer_calls_per_step'), d.get('peer_fused_per_step'), d.get('bce_in_range'))
PY
tail -3 gpurun_out/bench_8gpu_final.err
