cd $GRAFT_REPO_ROOT
timeout 600 python tools/gpu/diag_group.py 2>&1 | grep -v Warning | tail -120
