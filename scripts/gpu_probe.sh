#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
echo diff; python scripts/diff_probe.py 2>&1 | tail -5
echo tile; ALNR_NO_DIFF=1 python scripts/diff_probe.py 2>&1 | tail -5
