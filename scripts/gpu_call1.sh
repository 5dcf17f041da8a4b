#!/bin/bash
# first GPU call of round 2: tests, MODE-1 repro, bench, memcheck
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > $O/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
ALNR_WARP_INT_MODE1=1 timeout 300 python scripts/mode1_repro.py > $O/mode1.log 2>&1; echo "rc=$?" >> $O/mode1.log
timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" >> $O/bench_n1.err
timeout 600 compute-sanitizer --tool memcheck --log-file $O/san_memcheck.log python scripts/sanitize_small.py > $O/san_memcheck.out 2>&1; echo "rc=$?" >> $O/san_memcheck.out
tail -3 $O/pytest_gpu.log; tail -20 $O/mode1.log; tail -c 600 $O/bench_n1.err; tail -5 $O/san_memcheck.out; tail -5 $O/san_memcheck.log
