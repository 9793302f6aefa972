#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python tools/kernel_profile.py > gpurun_out/kernel_time_cupti_sk.txt 2>&1; echo "cupti rc=$?"; sed -n 5,60p gpurun_out/kernel_time_cupti_sk.txt | cut -c1-150
