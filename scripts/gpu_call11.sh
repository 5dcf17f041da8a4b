#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
for s in '[[100000,12800]]' '[[100000,100000]]'; do
export PROBE_SHAPES="$s"
timeout 100 python scripts/diff_probe.py 2> gpurun_out/stamps.txt | tail -1
python - <<'PY'
import re
rows=[l for l in open('gpurun_out/stamps.txt') if l.startswith('band') and 'own block 1' in l]
# last pass only: take the final set of lines (bands increasing)
last=[]
for l in rows:
    b=int(l.split()[1].rstrip(':'))
    if last and b < last[-1][0]: last=[]
    last.append((b,l))
import statistics as st
ramp=[float(re.search(r'own block 1 -> 72: ([0-9.]+)',l).group(1)) for b,l in last]
hand=[float(re.search(r"own block 1: (-?[0-9.]+)",l).group(1)) for b,l in last]
lag=[float(re.search(r'8192: ([0-9.]+) ([0-9.]+) ([0-9.]+)',l).group(2)) for b,l in last]
print("bands sampled",len(last),"ramp (block 1 -> 72) median %.2f us, above's 72 -> own block 1 median %.2f us, lag at 72 median %.2f us"%(st.median(ramp),st.median(hand),st.median(lag)))
for b,l in last[::6]: print(l.split('|')[0][:110])
PY
done
