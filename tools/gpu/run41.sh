#!/bin/bash
cd /root/repo
timeout 120 python tools/gpu/probe_m64.py
