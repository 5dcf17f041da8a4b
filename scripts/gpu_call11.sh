#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
export PROBE_SHAPES='[[100000,12800]]'
timeout 100 python scripts/diff_probe.py 2>&1 | tail -32
