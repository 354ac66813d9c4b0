#!/bin/bash
# Round 2, GPU call Q (1 GPU): A/B of the chunk / occupancy variants of the single-pass x-update
O=gpurun_out/r02_q; mkdir -p $O
for v in 0 1 2 3 4; do
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --opt xup_variant=$v > $O/bench_x$v.json 2> $O/bench_x$v.err; echo "x$v rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_q/bench_x*.json')):
    d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
    print(f.split('/')[-1], 'c2 %.4f xup %.4f | c3 %.4f xup %.4f | c4 %.4f xup %.4f' % (d['ms_per_step'], d['phase_ms_per_step']['xupdate'], d['workloads']['c3']['ms_per_step'], d['workloads']['c3']['phase_ms_per_step']['xupdate'], d['workloads']['c4']['ms_per_step'], d['workloads']['c4']['phase_ms_per_step']['xupdate']))
PY
