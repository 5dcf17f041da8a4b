#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
timeout 1200 python -m pytest tests/test_warp.py tests/test_heuristic.py tests/test_pwm.py tests/test_scores.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_warp.log 2>&1; echo "pytest rc=$?" >> $O/pytest_warp.log
tail -4 $O/pytest_warp.log
ALNR_WARP_INT_MODE1=1 timeout 900 python -m pytest tests/test_warp.py tests/test_scores.py tests/test_gpu_parity.py tests/test_pwm.py -m gpu -x -q > $O/pytest_mode1.log 2>&1; echo "pytest rc=$?" >> $O/pytest_mode1.log
tail -3 $O/pytest_mode1.log
python scripts/bench_configs.py c5big n1 n2 > $O/side_r02.jsonl 2> $O/side_r02.err; cat $O/side_r02.jsonl | cut -c1-250
