#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -15 $O/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_all_noalloc.log 2>&1; echo "all rc=$?"; tail -6 $O/pytest_all_noalloc.log
