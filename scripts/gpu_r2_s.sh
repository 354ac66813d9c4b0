#!/bin/bash
# Round 2, GPU call S (1 GPU): full single-GPU suite + bench of the library with the step tail on (mode 1)
O=gpurun_out/r02_s; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q --deselect tests/test_gpu_multi.py > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_s/bench.json').read().split('\n') if l.startswith('{')][0])
w=d['workloads']
print('c2 %.4f e2e %.4f | c3 %.4f e2e %.4f | c4 %.4f e2e %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], w['c3']['ms_per_step'], w['c3']['e2e']['ms_per_step'], w['c4']['ms_per_step'], w['c4']['e2e']['ms_per_step']))
print(d['phase_ms_per_step'], d['roofline'])
PY
