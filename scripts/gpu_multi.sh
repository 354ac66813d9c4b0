#!/bin/bash
# usage: gpu_multi.sh <tag> <ngpus> [pytest 0|1]
TAG=$1; N=$2; PT=${3:-1}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
if [ "$PT" = "1" ]; then
  timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -15 $O/pytest_multi.log
fi
for n in $N; do
  if [ "$n" = "1" ]; then
    timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > $O/bench_n$n.json 2> $O/bench_n$n.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 6 --warmup 3 --no-cpu-baseline > $O/bench_n$n.json 2> $O/bench_n$n.err
  fi
  echo "bench n=$n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_n$n.json").read().strip().split("\n")[-1]); print("n=$n %.3f G/s ms/step %.3f e2e %.3f G/s prox ms %.4f" % (d["value"]/1e9, d["ms_per_step"], d["e2e"]["value"]/1e9, d["roofline"]["avg_launch_ms"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()}, d["config"]["elements"])
except Exception as e: print("n=$n failed", e); print(open("$O/bench_n$n.err").read()[-1500:])
PY
done
