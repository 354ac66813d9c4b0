#!/bin/bash
TAG=$1; N=${2:-8}; O=gpurun_out/$TAG; mkdir -p $O
nproc > $O/nproc.txt
for wl in ${3:-c2 c4 c5}; do
if [ $wl = c5 ]; then ARGS="--workload c2 --nx 2048"; else ARGS="--workload $wl"; fi   # c5 = SURVEY C5 strong scaling (16.8 M triangles at every N)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N $ARGS --steps 6 --warmup 3 --no-cpu-baseline > $O/bench_${wl}_n$N.json 2> $O/bench_${wl}_n$N.err
python - <<PY
import json
try:
    d=json.loads(open("$O/bench_${wl}_n$N.json").read().strip().split("\n")[-1]); print("$wl n=$N %.3f G/s ms/step %.3f e2e %.3f G/s prox ms %.4f" % (d["value"]/1e9, d["ms_per_step"], d["e2e"]["value"]/1e9, d["roofline"]["avg_launch_ms"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()}, d["config"]["elements"], "nproc", open("$O/nproc.txt").read().strip())
except Exception as e: print("$wl n=$N failed", e); print(open("$O/bench_${wl}_n$N.err").read()[-1500:])
PY
done
