#!/bin/bash
# usage: gpu_opt.sh <tag> "<opt strings, one per run; '-' = none>" "<workloads>"
TAG=$1; OPTS=$2; WLS=${3:-"c2 c4"}; O=gpurun_out/$TAG; mkdir -p $O
for wl in $WLS; do for o in $OPTS; do
  if [ "$o" = "-" ]; then A=""; else A="--opt $o"; fi
  timeout 400 python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline $A > $O/${wl}_$o.json 2> $O/${wl}_$o.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${wl}_$o.json")); print("$wl $o %.3f G/s ms/step %.3f prox ms %.4f frac %.3f" % (d["value"]/1e9, d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()})
except Exception as e: print("$wl $o failed", e); print(open("$O/${wl}_$o.err").read()[-800:])
PY
done; done
