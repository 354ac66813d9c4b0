#!/bin/bash
TAG=$1; N=${2:-2}; STEPS=${3:-6}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps $STEPS --warmup 3 --no-cpu-baseline > $O/bench_n$N.json 2> $O/bench_n$N.err
python - <<PY
import json
d=json.loads(open("$O/bench_n$N.json").read().strip().split("\n")[-1]); print("n=$N %.3f G/s ms/step %.3f" % (d["value"]/1e9, d["ms_per_step"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()}, "step_ms", d["step_ms_rank0"])
PY
