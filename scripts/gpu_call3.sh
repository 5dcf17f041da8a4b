#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 1200 python -m pytest tests/test_wave.py tests/test_multi.py -m gpu -x -q > $O/pytest_ckpt.log 2>&1; echo "pytest rc=$?" >> $O/pytest_ckpt.log
tail -25 $O/pytest_ckpt.log
