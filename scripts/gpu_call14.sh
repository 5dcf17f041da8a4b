#!/bin/bash
# evidence for DESIGN 4.4: long-pair fills, both kernels, and long local pairs on both paths
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/long_pairs.txt
{
echo "# scripts/diff_probe.py: fill time of ONE long global DNA pair (+5/-4, del = ext = 8), device-resident, 3 passes"
echo "## difference-recurrence kernel (default)"
timeout 200 python scripts/diff_probe.py 2>&1 | tail -5
echo "## tile pipeline (ALNR_NO_DIFF=1)"
ALNR_NO_DIFF=1 timeout 200 python scripts/diff_probe.py 2>&1 | tail -5
echo
echo "# scripts/stateful_probe.py: ONE long LOCAL pair, del = 11, ext = 2 (stateful penalty), whole call;"
echo "# 'static' = the same pair with del = ext = 11; 'one warp' = the round-1 path (ALNR_NO_WAVE_STATEFUL=1)"
timeout 400 python scripts/stateful_probe.py 5000 20000 50000 100000 2>&1 | tail -16
} > $O 2>&1
cat $O
