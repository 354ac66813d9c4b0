#!/bin/bash
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
for wl in c2 c4 c3; do for mode in persist nopersist; do
  if [ "$mode" = "nopersist" ]; then export MMADMM_NO_L2PERSIST=1; else unset MMADMM_NO_L2PERSIST; fi
  timeout 400 python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline > $O/${wl}_$mode.json 2> $O/${wl}_$mode.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${wl}_$mode.json")); print("$wl $mode %.3f G/s ms/step %.3f prox ms %.4f frac %.3f" % (d["value"]/1e9, d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()})
except Exception as e: print("$wl $mode failed", e); print(open("$O/${wl}_$mode.err").read()[-800:])
PY
done; done
