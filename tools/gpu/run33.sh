#!/bin/bash
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -q -m gpu > gpurun_out/t_gpu_all4.log 2>&1
echo "gpu tests rc=$?"; tail -6 gpurun_out/t_gpu_all4.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_final2.log 2>gpurun_out/bench_final2.err
echo "bench rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/bench_final2.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['e2e'], d['gpu_launches'], d.get('bce_in_range'), d.get('peak_mem_gib'), d['roofline']['frac'], d['roofline']['hbm']['frac'], d.get('second_loop'), d.get('cpu_baseline'))
PY
tail -3 gpurun_out/bench_final2.err
