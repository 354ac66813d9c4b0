#!/bin/bash
# Round 2, GPU call F (2 GPUs): z/u slot-interleaved layout -- parity tests + bench; N=2 per-step timings
O=gpurun_out/r02_f; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py::test_a_stalled_rank_does_not_kill_its_peers > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 5 --per-step --no-extra-workloads --no-parity > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
def load(f):
    return json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
d=load('gpurun_out/r02_f/bench.json')
def show(r,ind=''):
    print(ind,'dev %.4f e2e %.4f strict %.4f frac %.3f' % (r['ms_per_step'], r['e2e']['ms_per_step'], r['e2e']['strict']['ms_per_step'], r['roofline']['frac']), {k:round(v,4) for k,v in r['phase_ms_per_step'].items()})
show(d)
for k,v in d['workloads'].items(): show(v,k)
d=load('gpurun_out/r02_f/bench_n2.json'); show(d,'N=2'); print(d['step_ms_rank0'])
PY
