#!/bin/bash
# full bench.py record on N GPUs of one box (N baked in by sed: scripts/gpu_bench_n.sh -> gpu_bench_nN.sh)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
N=1
if [ "$N" = "1" ]; then
  timeout 1200 python bench.py --gpus 1 --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench rc=$?"
else
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench rc=$?"
fi
tail -c 600 $O/bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','value_compact','ms_per_step','n_gpus')}, 'roof', d['roofline']['frac'], d['roofline']['traffic'])
print('e2e',d['e2e']['value'],d['e2e']['s_per_step'],'strings',d['e2e_strings'] and d['e2e_strings']['value'])
print('strong',d['strong_scaling'])
for k,v in d['configs'].items(): print(k, round(v['gcups'],2), v['parity'], v.get('fitness_sha256_16'), v.get('roofline',{}).get('frac'))
print('cpu', d.get('cpu_baseline'))
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cut -c1-300 $O/bench_ref.json
