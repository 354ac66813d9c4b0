#!/bin/bash
# Round 2, GPU call J (1 GPU): incidence-ordered z -- whole single-GPU suite, bench, A/B of the x-update variants
O=gpurun_out/r02_j; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
for v in 1 2 3; do
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --opt xup_variant=$v > $O/bench_x$v.json 2> $O/bench_x$v.err; echo "bench x$v rc=$?"
done
python - <<'PY'
import json,glob
def show(r,ind=''):
    print(ind,'dev %.4f e2e %.4f strict %.4f frac %.3f' % (r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step'], r['roofline']['frac']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()})
for f in sorted(glob.glob('gpurun_out/r02_j/bench*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
    except Exception as e:
        print(f, e); continue
    show(d, f.split('/')[-1])
    for k,v in d['workloads'].items(): show(v,'   '+k)
PY
