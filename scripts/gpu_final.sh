#!/bin/bash
# usage: gpu_final.sh <tag> -- last check of a round: full GPU suite, smoke(), the three bench lines and the reference arm
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke.log
python bench.py --steps 10 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench c2 rc=$?"
python bench.py --workload c3 --steps 10 --warmup 3 > $O/bench_c3.json 2> $O/bench_c3.err; echo "bench c3 rc=$?"
python bench.py --workload c4 --steps 6 --warmup 3 > $O/bench_c4.json 2> $O/bench_c4.err; echo "bench c4 rc=$?"
python - <<PY
import json
for wl in ("c2","c3","c4"):
    d=json.loads(open("$O/bench_%s.json" % wl).read().strip().split("\n")[-1])
    print(wl, "%.3f ms/step %.3f G/s e2e %.3f G/s prox %.4f ms frac %.3f" % (d["ms_per_step"], d["value"]/1e9, d["e2e"]["value"]/1e9, d["roofline"]["avg_launch_ms"], d["roofline"]["frac"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()}, d["clocks"])
PY
