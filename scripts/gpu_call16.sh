#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
export PROBE_SHAPES='[[100000,12800],[100000,60000],[100000,100000],[140000,140000]]'
for g in 0 1184 888 740 592 444 296; do
  echo "grid $g"
  if [ $g = 0 ]; then timeout 120 python scripts/diff_probe.py 2>&1 | tail -4; else ALNR_DIFF_GRID=$g timeout 120 python scripts/diff_probe.py 2>&1 | tail -4; fi
done
