cd $GRAFT_REPO_ROOT
timeout 900 python tools/gpu/diag_streams.py 2>&1 | grep -v Warning | tail -12
