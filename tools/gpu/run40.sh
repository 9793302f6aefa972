#!/bin/bash
cd /root/repo
timeout 300 python tools/flat_trace.py 48 48 5 64
timeout 300 python tools/flat_trace.py 48 48 3 64
timeout 300 python tools/flat_trace.py 96 96 3 64
