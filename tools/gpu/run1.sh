set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L | head -3; nproc
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" 
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu_all.log 2>&1; echo "gpu tests rc=$?"
tail -5 gpurun_out/t_gpu_all.log
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_start.log 2>&1; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_start.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 --cpu-seconds 50 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?"; tail -c 600 gpurun_out/bench_ref.log
GANB200_STYLE_BANK=1 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_bank1.log 2>&1; echo "bank1 rc=$?"; tail -c 400 gpurun_out/bench_bank1.log | cut -c1-300
GANB200_STYLE_BANK=2 timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu --second-loop '' > gpurun_out/bench_bank2.log 2>&1; echo "bank2 rc=$?"; tail -c 400 gpurun_out/bench_bank2.log | cut -c1-300
timeout 600 python tools/kernel_profile.py 256 ganR2 > gpurun_out/cupti_start.txt 2>&1; echo "cupti rc=$?"; head -30 gpurun_out/cupti_start.txt
