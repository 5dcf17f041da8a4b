#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
python scripts/prof_c5.py 2000 100 > $O/prof_c5_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fill_warp_kernel -s 0 -c 1 -f -o $O/c5_r02 python scripts/prof_c5.py 2000 100 > $O/ncu_c5.log 2>&1
echo "rc=$?"; cat $O/prof_c5_plain.log; tail -3 $O/ncu_c5.log
python scripts/prof_c5.py 2000 100 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c5_launches_r02.csv python scripts/prof_c5.py 2000 100 > /dev/null 2>&1
