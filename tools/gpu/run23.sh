#!/bin/bash
set -x
cd /root/repo
mkdir -p gpurun_out
timeout 900 python tools/gpu/diag_det2.py > gpurun_out/diag_det2.log 2>&1
echo "rc=$?"; tail -40 gpurun_out/diag_det2.log
