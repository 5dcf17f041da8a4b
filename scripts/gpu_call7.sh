#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest_all7.log 2>&1; echo "pytest rc=$?"; tail -16 $O/pytest_all7.log
python scripts/c4_after_c3.py 2>&1 | head -1
