#!/bin/bash
# Round 2, GPU call AQ (2 GPUs): multi-GPU tests (world 2) and the N=2 bench line + reference arm of the final library
O=gpurun_out/r02_aq; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -q --durations=6 > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -4 $O/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_aq/bench_n2.json').read().split('\n') if l.startswith('{')][0])
print('n2 value %.4g ms %.4f e2e %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value']), d.get('parity'))
for k,v in d.get('workloads',{}).items(): print(k, '%.4g' % v['value'], v['ms_per_step'])
PY
