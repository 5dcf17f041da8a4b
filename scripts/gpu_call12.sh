#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 600 python -m pytest tests/test_wave.py -m gpu -x -q -k "stateful" > $O/pytest_wave12.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest_wave12.log | cut -c1-220
