#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
export PROBE_SHAPES='[[30000,30000],[50000,50000],[70000,70000],[85000,85000],[100000,60000],[60000,100000],[100000,100000],[140000,140000]]'
echo DIFF; timeout 200 python scripts/diff_probe.py 2>&1 | tail -8
echo TILE; ALNR_NO_DIFF=1 timeout 200 python scripts/diff_probe.py 2>&1 | tail -8
