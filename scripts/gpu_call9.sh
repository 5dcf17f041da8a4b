#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 300 python -m pytest tests/test_wave.py -m gpu -x -q -k "diff or wavefront or extreme" > $O/pytest_wave9.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_wave9.log
timeout 120 python scripts/diff_probe.py 2>&1 | tail -5
echo TILE; ALNR_NO_DIFF=1 timeout 120 python scripts/diff_probe.py 2>&1 | tail -5
