#!/bin/bash
# GPU session B: parity tests, prox-variant sweep on C2 and C4, memcheck of the smoke run.
mkdir -p gpurun_out/b
O=gpurun_out/b
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest.log
tail -5 $O/pytest.log
for v in 0 1 2 3 4; do
  timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --prox-variant $v > $O/c2_v$v.json 2> $O/c2_v$v.err
  python - <<PY
import json
try:
    d=json.load(open("$O/c2_v$v.json")); print("c2 v$v", d["value"]/1e9, "G/s ms/step", d["ms_per_step"], "prox ms", d["roofline"]["avg_launch_ms"], "frac", d["roofline"]["frac"], d["phase_ms_per_step"])
except Exception as e: print("c2 v$v failed", e)
PY
done
for v in 0 1; do
  timeout 400 python bench.py --workload c4 --steps 6 --warmup 3 --no-cpu-baseline --prox-variant $v > $O/c4_v$v.json 2> $O/c4_v$v.err
  python - <<PY
import json
try:
    d=json.load(open("$O/c4_v$v.json")); print("c4 v$v", d["value"]/1e9, "G/s ms/step", d["ms_per_step"], "prox ms", d["roofline"]["avg_launch_ms"], "frac", d["roofline"]["frac"], d["phase_ms_per_step"])
except Exception as e: print("c4 v$v failed", e)
PY
done
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py smoke > $O/memcheck.log 2>&1; echo "memcheck rc=$?" | tee -a $O/memcheck.log
tail -3 $O/memcheck.log
