#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
for c in 0 1; do GANB200_CPAD64=$c timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_cpad$c.log 2>/dev/null; done
python - <<'PY'
import json
for f in ('bench_cpad0','bench_cpad1'):
    for l in open(f'gpurun_out/{f}.log'):
        if l.startswith('{'):
            d=json.loads(l); print(f, d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d.get('peak_mem_gib'), d['roofline']['frac'])
PY
