#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
for w in 2 3; do GANB200_BRANCH_STREAMS=$w timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_bs$w.log 2>gpurun_out/bench_bs$w.err; echo "width $w rc=$?"; tail -2 gpurun_out/bench_bs$w.err | cut -c1-300; done
python - <<'PY'
import json
for f in ('bench_bs2','bench_bs3'):
    try:
        for l in open(f'gpurun_out/{f}.log'):
            if l.startswith('{'):
                d=json.loads(l); print(f, d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d.get('peak_mem_gib'), d.get('losses'))
    except Exception as e: print(f, e)
PY
