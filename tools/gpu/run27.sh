#!/bin/bash
# full GPU suite after the deterministic reductions / peer exchange / BN cap; ncu --set full of the 48-channel conv
# kernels (fprop, dgrad, wgrad), the modulated-conv backward (fold_scale_reduce) and the pack kernel
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -q -m gpu > gpurun_out/t_gpu_all3.log 2>&1
echo "gpu tests rc=$?"; tail -6 gpurun_out/t_gpu_all3.log
for op in fprop dgrad wgrad; do
  python tools/one_conv.py 48 48 5 64 1 $op && ncu --set full --clock-control none --import-source on -k regex:'tc_' -s 2 -c 2 -f -o gpurun_out/r02_c48_$op python tools/one_conv.py 48 48 5 64 1 $op > gpurun_out/ncu_c48_$op.log 2>&1; echo "ncu c48 $op rc=$?"
done
python tools/one_modbwd.py && ncu --set full --clock-control none --import-source on -k regex:'fold_scale_reduce|pack_act' -s 3 -c 3 -f -o gpurun_out/r02_modbwd_hbm python tools/one_modbwd.py > gpurun_out/ncu_modbwd.log 2>&1; echo "ncu modbwd rc=$?"
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_bn8192.log 2>gpurun_out/bench_bn8192.err
echo "bench rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/bench_bn8192.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d.get('losses'))
PY
