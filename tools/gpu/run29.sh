#!/bin/bash
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 120 python tools/gpu/probe_shift.py > gpurun_out/probe_shift.log 2>&1; echo "probe rc=$?"; cat gpurun_out/probe_shift.log
timeout 900 python tools/kernel_profile.py > gpurun_out/kernel_time_cupti_det.txt 2>&1; echo "cupti rc=$?"; head -30 gpurun_out/kernel_time_cupti_det.txt | cut -c1-180
