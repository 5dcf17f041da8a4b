#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_compact.py tests/test_scores.py tests/test_full_size.py -m gpu -x -q > $O/pytest_strip.log 2>&1; echo "pytest rc=$?" >> $O/pytest_strip.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-configs --no-e2e-strings --no-cpu-baseline > $O/bench_v2.json 2> $O/bench_v2.err; echo "bench rc=$?" >> $O/bench_v2.err
tail -5 $O/pytest_strip.log; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_v2.json'))
print({k:d[k] for k in ('value','value_compact','ms_per_step')}, d['roofline']['fill_ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['config']['parity_spot_check'])
PY
