#!/bin/bash
# Round 2, GPU call P (1 GPU): single-pass primal in the streaming x-update -- single-GPU suite, bench
O=gpurun_out/r02_p; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_p/bench.json').read().split('\n') if l.startswith('{')][0])
def show(r,ind=''):
    print(ind,'dev %.4f e2e %.4f strict %.4f frac %.3f' % (r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step'], r['roofline']['frac']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()})
show(d)
for k,v in d['workloads'].items(): show(v,k)
PY
