#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 120 python scripts/prof_c4.py 100000 640 > $O/prof_c4_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fill_diff_kernel -s 1 -c 1 -f -o $O/diff_ramp python scripts/prof_c4.py 100000 640 > $O/ncu_c4.log 2>&1
echo "ncu rc=$?"; cat $O/prof_c4_plain.log
