#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_all5.log 2>&1; echo "pytest rc=$?" >> $O/pytest_all5.log
tail -4 $O/pytest_all5.log
python scripts/bench_configs.py c5big > $O/side5.jsonl 2>&1; cut -c1-200 $O/side5.jsonl
ALNR_HEUR_ONE_GROUP=1 python scripts/bench_configs.py c5big > $O/side5b.jsonl 2>&1; cut -c1-200 $O/side5b.jsonl
timeout 600 python bench.py --steps 5 --warmup 3 --no-configs --no-e2e-strings --no-cpu-baseline > $O/bench_v3.json 2> $O/bench_v3.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_v3.json'))
print({k:d[k] for k in ('value','value_compact','ms_per_step')}, d['roofline']['fill_ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['config']['parity_spot_check'])
PY
