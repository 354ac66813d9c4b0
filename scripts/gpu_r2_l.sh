#!/bin/bash
# Round 2, GPU call L (2 GPUs): whole GPU suite on the current library (world-2 multi-GPU cases included), N=2 bench (all workloads)
O=gpurun_out/r02_l; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q --durations=15 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest_gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json,glob
def show(r,ind=''):
    print(ind,'value %.4g dev %.4f e2e %.4f strict %.4f' % (r['value'], r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()}, 'launches', r['gpu_launches'])
for f in sorted(glob.glob('gpurun_out/r02_l/bench*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
    except Exception as e:
        print(f, e); continue
    show(d, f.split('/')[-1])
    if 'parity' in d: print('   parity', d['parity']['max_rel_dx'], d['parity']['max_rel_Ih'], d['parity']['iters_equal'])
    for k,v in d.get('workloads',{}).items(): show(v,'   '+k)
PY
