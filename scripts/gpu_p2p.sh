#!/bin/bash
# usage: gpu_p2p.sh <tag> "<ngpus list>" [pytest 0|1] -- multi-GPU tests, then bench with peer-memory exchange and with NCCL
TAG=$1; NS=$2; PT=${3:-1}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
if [ "$PT" = "1" ]; then
  timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q > $O/pytest_multi.log 2>&1; echo "pytest multi (p2p) rc=$?"; tail -4 $O/pytest_multi.log
  MMADMM_NO_P2P=1 timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q -k "world0 or 2-" > $O/pytest_multi_nccl.log 2>&1; echo "pytest multi (nccl) rc=$?"; tail -2 $O/pytest_multi_nccl.log
fi
for n in $NS; do for mode in p2p nccl; do
  if [ "$mode" = "nccl" ]; then export MMADMM_NO_P2P=1; else unset MMADMM_NO_P2P; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 6 --warmup 3 --no-cpu-baseline $EXTRA > $O/bench_n${n}_$mode.json 2> $O/bench_n${n}_$mode.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_n${n}_$mode.json").read().strip().split("\n")[-1]); print("n=$n $mode %.3f G/s ms/step %.3f e2e %.3f G/s prox ms %.4f" % (d["value"]/1e9, d["ms_per_step"], d["e2e"]["value"]/1e9, d["roofline"]["avg_launch_ms"]), {k: round(x,4) for k,x in d["phase_ms_per_step"].items()}, "p2p" if "NVLink" in d["config"]["parallelism"] else "nccl")
except Exception as e: print("n=$n $mode failed", e); print(open("$O/bench_n${n}_$mode.err").read()[-1500:])
PY
done; done
