#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/prot_launches.csv python scripts/prof_protein.py 20000 > $O/ncu_protl.log 2>&1
echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fill_warp_kernel -s 0 -c 1 -f -o $O/prot_r02 python scripts/prof_protein.py 20000 > $O/ncu_prot.log 2>&1
echo "ncu rc=$?"
