#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_compact.py tests/test_full_size.py tests/test_scores.py tests/test_multi.py -m gpu -x -q > $O/pytest6.log 2>&1; echo "pytest rc=$?" >> $O/pytest6.log; tail -3 $O/pytest6.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-configs --no-e2e-strings --no-cpu-baseline > $O/bench_v4.json 2> $O/bench_v4.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_v4.json'))
print({k:d[k] for k in ('value','value_compact','ms_per_step')}, d['roofline']['fill_ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['config']['parity_spot_check'])
PY
python scripts/prof_c3.py 1600000 > $O/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fill_strip_kernel -s 4 -c 1 -f -o $O/strip_r02b python scripts/prof_c3.py 1600000 > $O/ncu_strip.log 2>&1
echo "ncu rc=$?"; cat $O/prof_plain.log
