#!/bin/bash
# Round 2, GPU call K (8 GPUs): multi-GPU tests at world 2/4/8, D2H study, bench ladder N=8 (+ N=4, N=2 c2 only)
O=gpurun_out/r02_k; mkdir -p $O
nvidia-smi -L > $O/gpus.txt; nproc > $O/nproc.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 1800 python -m pytest tests/test_gpu_multi.py -m gpu -q --durations=10 > $O/pytest_multi_8gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_multi_8gpu.log
for n in 1 2 4 8; do timeout 300 $TR --nproc-per-node $n --master-port 2951$n scripts/d2h_bw.py 2>/dev/null | grep '^{' > $O/d2h_n$n.json; cat $O/d2h_n$n.json; done
timeout 1200 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench n8 rc=$?"
timeout 600 $TR --nproc-per-node 4 --master-port 29522 bench.py --gpus 4 --steps 20 --warmup 5 --no-extra-workloads > $O/bench_n4.json 2> $O/bench_n4.err; echo "bench n4 rc=$?"
timeout 600 $TR --nproc-per-node 2 --master-port 29523 bench.py --gpus 2 --steps 20 --warmup 5 --no-extra-workloads > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --nx 2048 --steps 20 --warmup 5 --no-extra-workloads --no-parity > $O/bench_c5_strong_n8.json 2> $O/bench_c5_strong_n8.err; echo "bench c5 strong n8 rc=$?"
python - <<'PY'
import json,glob
def show(r,ind=''):
    print(ind,'value %.4g dev %.4f e2e %.4f strict %.4f' % (r['value'], r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()}, 'launches', r['gpu_launches'])
for f in sorted(glob.glob('gpurun_out/r02_k/bench*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
    except Exception as e:
        print(f, e); continue
    show(d, f.split('/')[-1])
    if 'parity' in d: print('   parity', d['parity']['max_rel_dx'], d['parity']['max_rel_Ih'], d['parity']['iters_equal'])
    for k,v in d.get('workloads',{}).items(): show(v,'   '+k)
PY
