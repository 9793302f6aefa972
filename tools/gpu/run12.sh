cd $GRAFT_REPO_ROOT
timeout 600 python tools/scatter_bench.py 256 2>&1 | tail -20
timeout 600 python -m pytest tests/test_golden_final_pair.py -q -m gpu -x 2>&1 | grep -E "^E |bad|assert" | head -12
