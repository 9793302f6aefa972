#!/bin/bash
# deterministic reductions: kernel parity, graph-vs-eager distance, bench
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py tests/test_zz_no_alloc.py -q -m gpu -x > gpurun_out/t_det_kernels.log 2>&1
echo "kernels rc=$?"; tail -5 gpurun_out/t_det_kernels.log
timeout 1500 python -m pytest tests/test_graphed_step.py -q -m gpu -s -k "final_pair or mlp_pair" > gpurun_out/t_det_graph.log 2>&1
echo "graph rc=$?"; grep -a "eager vs eager\|passed\|failed\|Error" gpurun_out/t_det_graph.log | tail -8
timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_det.log 2>gpurun_out/bench_det.err
echo "bench rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/bench_det.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d.get('losses'))
PY
tail -3 gpurun_out/bench_det.err
