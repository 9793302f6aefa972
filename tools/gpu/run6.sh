cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_dist_nccl.py -q -m gpu -x 2>&1 | grep -v Warning | grep -E "^E |assert|Error" | head -30
