#!/bin/bash
# usage: gpu_profile.sh <tag>   -- bench line + ncu launch list + ncu full capture of the prox kernel (C2), bench line for C4
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
python bench.py --steps 10 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench c2 rc=$?"
python bench.py --workload c4 --steps 6 --warmup 3 > $O/bench_c4.json 2> $O/bench_c4.err; echo "bench c4 rc=$?"
python bench.py --impl reference --steps 4 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv $CMD > $O/ncu_list.log 2>&1
echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_prox -s 8 -c 1 -f -o $O/prof_c2 $CMD > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu --set full --clock-control none -k regex:k_xupdate -s 12 -c 2 -f -o $O/prof_xup $CMD > $O/ncu_xup.log 2>&1
echo "ncu xup rc=$?"
