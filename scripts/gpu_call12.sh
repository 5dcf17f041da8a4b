#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 600 python -m pytest tests/test_wave.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_wave12.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest_wave12.log | cut -c1-220
timeout 300 python scripts/stateful_probe.py 20000 100000 2>&1 | tail -8
