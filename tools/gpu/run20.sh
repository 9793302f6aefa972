cd $GRAFT_REPO_ROOT
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -6
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_final.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_final.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['e2e'], d['gpu_launches'], d.get('bce_in_range'), d['peak_mem_gib'], d['roofline']['frac'], d['roofline']['hbm']['frac'], d.get('second_loop'), d.get('cpu_baseline'))"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_final_ref.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/bench_final_ref.log | cut -c1-300
