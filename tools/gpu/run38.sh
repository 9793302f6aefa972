#!/bin/bash
cd /root/repo
for cfg in "2 2" "2 3" "1 2" "1 3" "1 4"; do set -- $cfg; echo "== mt=$1 a_stages=$2"; GANB200_FLAT_MT=$1 GANB200_FLAT_ASTAGES=$2 timeout 300 python tools/flat_bench.py 2>&1 | sed -n 3,9p; done
