#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py tests/test_golden_final_pair.py -q -m gpu -x 2>&1 | tail -3
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_demod.log 2>/dev/null
python - <<'PY'
import json
for l in open('gpurun_out/bench_demod.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d.get('losses'))
PY
