#!/bin/bash
# Round 2, GPU call AX (8 GPUs): N=8 bench line of the final library (parity record inside), nothing else (8x charged)
O=gpurun_out/r02_ax; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench n8 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_ax/bench_n8.json').read().split('\n') if l.startswith('{')][0])
print('n8 value %.4g ms %.4f e2e %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value']), {k: d['parity'][k] for k in ('max_rel_dx','max_rel_Ih','iters_equal','ok')}, d['phase_ms_per_step'])
for k,v in d.get('workloads',{}).items(): print(k, '%.4g' % v['value'], v['ms_per_step'])
PY
