#!/bin/bash
O=gpurun_out/r02_g; mkdir -p $O
for rep in 1 2 3; do for per in 2 0; do
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra-workloads --per-step --clock-period-ms $per > $O/c2_p${per}_$rep.json 2> $O/c2_p${per}_$rep.err
done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_g/*.json')):
    d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
    print(f.split('/')[-1], 'dev %.4f e2e %.4f strict %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['strict']['ms_per_step']), 'e2e steps', d['e2e']['step_ms_rank0'])
PY
