#!/bin/bash
# Round 2, GPU call AP (1 GPU): the final library once more -- whole GPU suite, smoke(), default bench line, Method 2 line
O=gpurun_out/r02_ap; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q -x --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
timeout 600 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 $O/smoke.log
python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench (no flags) rc=$?"
python bench.py --method 2 --steps 5 --warmup 1 --no-cpu-baseline > $O/bench_m2.json 2> $O/bench_m2.err; echo "bench method 2 rc=$?"
python - <<'PY'
import json
for f in ('bench_default','bench_m2'):
    d=json.loads([l for l in open('gpurun_out/r02_ap/%s.json'%f).read().split('\n') if l.startswith('{')][0])
    print(f, d['value'], d['unit'], d['ms_per_step'], d.get('e2e',{}).get('value'), d.get('roofline',{}).get('frac'))
PY
