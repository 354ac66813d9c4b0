#!/bin/bash
# Round 2, GPU call O (8 GPUs): final library -- N=8 default line (c2 + c3 + c4 + parity), C5 strong scaling at N=8
O=gpurun_out/r02_o; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 1200 $TR --nproc-per-node 8 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench n8 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29542 bench.py --gpus 8 --nx 2048 --steps 20 --warmup 5 --no-extra-workloads --no-parity > $O/bench_c5_strong_n8.json 2> $O/bench_c5_strong_n8.err; echo "bench c5 strong n8 rc=$?"
timeout 600 $TR --nproc-per-node 4 --master-port 29543 bench.py --gpus 4 --steps 20 --warmup 5 --no-extra-workloads --no-parity > $O/bench_n4.json 2> $O/bench_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json,glob
def show(r,ind=''):
    print(ind,'value %.4g dev %.4f e2e %.4f strict %.4f' % (r['value'], r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()}, 'launches', r['gpu_launches'])
for f in sorted(glob.glob('gpurun_out/r02_o/bench*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
    except Exception as e:
        print(f, e); continue
    show(d, f.split('/')[-1])
    if 'parity' in d: print('   parity', d['parity']['max_rel_dx'], d['parity']['max_rel_Ih'], d['parity']['iters_equal'])
    for k,v in d.get('workloads',{}).items(): show(v,'   '+k)
PY
