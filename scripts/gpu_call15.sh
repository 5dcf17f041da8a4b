#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --no-e2e-strings > $O/bench_try.json 2> $O/bench_try.err; echo "bench rc=$?"; tail -c 400 $O/bench_try.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_try.json').read().strip().splitlines()[-1])
for k,v in d['configs'].items(): print(k, round(v['gcups'],2), round(v.get('ms',0),3), v['parity'], v.get('cpu_port_gcups'))
PY
