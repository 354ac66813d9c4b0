#!/bin/bash
# Round 2, GPU call H (1 GPU): bench with the write-back set-up outside the step; ncu --set full of the 3D prox kernel and of the x-update
O=gpurun_out/r02_h; mkdir -p $O
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
CMD4="python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_prox_q -s 8 -c 1 -f -o $O/prof_c4 $CMD4 > $O/ncu_full_c4.log 2>&1; echo "ncu c4 rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra-workloads"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_xupdate -s 12 -c 1 -f -o $O/prof_xup $CMD > $O/ncu_xup.log 2>&1; echo "ncu xup rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_h/bench.json').read().split('\n') if l.startswith('{')][0])
def show(r,ind=''):
    print(ind,'dev %.4f e2e %.4f strict %.4f frac %.3f' % (r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step'], r['roofline']['frac']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()})
show(d)
for k,v in d['workloads'].items(): show(v,k)
PY
