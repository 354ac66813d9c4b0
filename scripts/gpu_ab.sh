#!/bin/bash
# usage: gpu_ab.sh <tag> -- gpu tests, then A/B of the isotropic grid path on c2/c3/c4
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest.log
for wl in c2 c3 c4; do for iso in 1 0; do
  timeout 400 python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline --iso-grid $iso > $O/${wl}_iso$iso.json 2> $O/${wl}_iso$iso.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${wl}_iso$iso.json")); print("$wl iso$iso %.3f G/s ms/step %.3f prox ms %.4f frac %.3f" % (d["value"]/1e9, d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()}, d["config"].get("monitor_grid_isotropic_fast_path"))
except Exception as e: print("$wl iso$iso failed", e); print(open("$O/${wl}_iso$iso.err").read()[-800:])
PY
done; done
