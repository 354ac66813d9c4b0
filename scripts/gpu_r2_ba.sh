#!/bin/bash
# Round 2, GPU call BA (1 GPU): 3D prox with the thread's own B rows kept in registers (prox_variant 3-5) against the default
O=gpurun_out/r02_ba; mkdir -p $O
for v in 0 3 4 5; do
  timeout 300 python bench.py --workload c4 --steps 10 --warmup 5 --no-cpu-baseline --opt prox_variant=$v > $O/bench_c4_v$v.json 2> $O/bench_c4_v$v.err; echo "v$v rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_ba/bench_c4_v*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
        print(f.split('/')[-1], 'ms %.4f prox %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['frac']), d.get('energy_last', ''), d['phase_ms_per_step'])
    except Exception as e: print(f, 'ERR', e)
PY
