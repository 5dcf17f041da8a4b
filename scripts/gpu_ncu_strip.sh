#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
python scripts/prof_c3.py 1600000 > $O/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fill_strip_kernel -s 4 -c 1 -f -o $O/strip_r02 python scripts/prof_c3.py 1600000 > $O/ncu_strip.log 2>&1
echo "rc=$?"; cat $O/prof_plain.log; tail -3 $O/ncu_strip.log
