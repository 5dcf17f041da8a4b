#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 200 python scripts/prof_protein.py 20000 > $O/prof_prot_plain.log 2>&1
timeout 200 python scripts/prof_protein.py 20000 11 11 >> $O/prof_prot_plain.log 2>&1
cat $O/prof_prot_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fill_warp_kernel -s 1 -c 1 -f -o $O/prot_r02 python scripts/prof_protein.py 20000 > $O/ncu_prot.log 2>&1
echo "ncu rc=$?"
