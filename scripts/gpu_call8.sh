#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 900 python -m pytest tests/test_wave.py -m gpu -x -q > $O/pytest_wave8.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest_wave8.log
timeout 120 python scripts/c4_after_c3.py 2>&1 | head -1
ALNR_NO_DIFF=1 timeout 120 python scripts/c4_after_c3.py 2>&1 | head -1
