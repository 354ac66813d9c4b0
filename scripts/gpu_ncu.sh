#!/bin/bash
# usage: gpu_ncu.sh <tag> <workload> <variant> [kernel regex]
TAG=$1; WL=$2; V=$3; K=${4:-k_prox}
O=gpurun_out/$TAG; mkdir -p $O
CMD="python bench.py --workload $WL --steps 3 --warmup 3 --no-cpu-baseline --prox-variant $V"
$CMD > $O/plain_$WL.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s 8 -c 2 -f -o $O/prof_$WL $CMD > $O/ncu_$WL.log 2>&1
echo "ncu rc=$?"; tail -3 $O/ncu_$WL.log
