#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 300 python -m pytest tests/test_wave.py -m gpu -x -q -k "diff or wavefront or extreme" 2>&1 | tail -2
export PROBE_SHAPES='[[128,100000],[12800,100000],[100000,12800],[100000,60000],[100000,100000],[140000,140000]]'
timeout 120 python scripts/diff_probe.py 2>&1 | tail -6
