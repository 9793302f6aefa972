#!/bin/bash
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/modbwd_bench.py > gpurun_out/modbwd_bench.log 2>&1; echo "rc=$?"; cat gpurun_out/modbwd_bench.log
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_style_layer.py -q -m gpu -x 2>&1 | tail -3
GANB200_BN_SMALL_MAX=4096 timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_bn4096.log 2>/dev/null
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_bn8192.log 2>/dev/null
python - <<'PY'
import json
for f in ('bench_bn4096','bench_bn8192'):
    for l in open(f'gpurun_out/{f}.log'):
        if l.startswith('{'):
            d=json.loads(l); print(f, d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'))
PY
