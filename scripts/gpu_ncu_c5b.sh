#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
python scripts/prof_c5.py 2000 100 > $O/prof_c5_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c5_launches_r02b.csv python scripts/prof_c5.py 2000 100 > /dev/null 2>&1
echo rc=$?
