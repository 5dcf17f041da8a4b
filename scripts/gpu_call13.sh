#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 120 python -m pytest tests/test_wave.py -m gpu -x -q -k stateful 2>&1 | tail -2
timeout 300 python scripts/stateful_probe.py 5000 20000 50000 100000 2>&1 | grep -v "one warp" | tail -12
