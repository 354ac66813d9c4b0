#!/bin/bash
# usage: gpu_sweep.sh <tag> <workload> "<variants>" [pytest: 0|1]
TAG=$1; WL=$2; VARS=$3; PT=${4:-1}
O=gpurun_out/$TAG; mkdir -p $O
if [ "$PT" = "1" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest.log
  tail -4 $O/pytest.log
fi
for v in $VARS; do
  timeout 400 python bench.py --workload $WL --steps 6 --warmup 3 --no-cpu-baseline --prox-variant $v > $O/${WL}_v$v.json 2> $O/${WL}_v$v.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${WL}_v$v.json")); print("$WL v$v %.3f G/s ms/step %.3f prox ms %.4f frac %.3f" % (d["value"]/1e9, d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()})
except Exception as e: print("$WL v$v failed", e); print(open("$O/${WL}_v$v.err").read()[-800:])
PY
done
