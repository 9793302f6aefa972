#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_golden_final_pair.py tests/test_graphed_step.py tests/test_style_layer.py -q -m gpu -x 2>&1 | tail -4
for f in 0 1; do GANB200_SK_BWD_FUSED=$f timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_skb$f.log 2>/dev/null; done
python - <<'PY'
import json
for f in ('bench_skb0','bench_skb1'):
    for l in open(f'gpurun_out/{f}.log'):
        if l.startswith('{'):
            d=json.loads(l); print(f, d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d.get('peak_mem_gib'), d.get('losses'))
PY
