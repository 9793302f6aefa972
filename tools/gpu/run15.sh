cd $GRAFT_REPO_ROOT
for m in 0 1; do GANB200_BRANCH_STREAMS=$m timeout 900 python bench.py --steps 6 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_streams$m.log 2>&1; echo "bench streams=$m rc=$?"; tail -1 gpurun_out/bench_streams$m.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['gpu_launches'], d.get('bce_in_range'), d['peak_mem_gib'], d['roofline']['frac'], d['losses'])" || tail -5 gpurun_out/bench_streams$m.log; done
timeout 1800 python -m pytest tests -q -m gpu -x 2>&1 | tail -6
