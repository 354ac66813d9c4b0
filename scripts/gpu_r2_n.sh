#!/bin/bash
# Round 2, GPU call N (1 GPU): 3D prox with the single-sweep pass 2 -- 3D parity tests, c4 bench; ncu --set full of k_tri_cluster
O=gpurun_out/r02_n; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_large.py tests/test_gpu_fullsize.py -m gpu -q -x -k "3D or 3d or c4 or 3-" > $O/pytest_3d.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_3d.log
timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err; echo "bench c4 rc=$?"
CMDM="python bench.py --method 2 --nx 512 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none -k regex:tri_cluster -c 2 -f -o $O/prof_tri $CMDM > $O/ncu_tri.log 2>&1; echo "ncu tri rc=$?"; tail -3 $O/ncu_tri.log
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_n/bench_c4.json').read().split('\n') if l.startswith('{')][0])
print('c4 dev %.4f e2e %.4f frac %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['frac']), d['phase_ms_per_step'])
PY
