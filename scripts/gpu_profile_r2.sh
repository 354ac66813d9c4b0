#!/bin/bash
# usage: gpu_profile_r2.sh <tag>  -- round-2 evidence run on ONE B200: GPU test suite, bench lines (default line = c2 + c3 + c4
# sub-records; reference arm; Method 2 line), ncu launch lists (c2, c4), ncu --set full captures of the prox kernels (2D on c2,
# 3D on c4), the x-update and the Method 2 cluster triangular solve.  Every ncu pass runs only after its command exited 0 without ncu.
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q -s --durations=20 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench (default line) rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
python bench.py --nx 2048 --steps 10 --warmup 5 --no-cpu-baseline --no-extra-workloads > $O/bench_c5_n1.json 2> $O/bench_c5_n1.err; echo "bench c5 (nx=2048) rc=$?"
python bench.py --method 2 --steps 5 --warmup 1 > $O/bench_m2.json 2> $O/bench_m2.err; echo "bench method 2 rc=$?"
python bench.py --workload c4 --steps 10 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err; echo "bench c4 rc=$?"
python bench.py --workload c3 --steps 10 --warmup 5 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo "bench c3 rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra-workloads"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv $CMD > $O/ncu_list.log 2>&1
echo "ncu list c2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_prox_p -s 8 -c 1 -f -o $O/prof_c2 $CMD > $O/ncu_full.log 2>&1
echo "ncu full c2 rc=$?"
ncu --set full --clock-control none -k regex:k_xupdate -s 12 -c 2 -f -o $O/prof_xup $CMD > $O/ncu_xup.log 2>&1
echo "ncu xup rc=$?"
CMD4="python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline"
$CMD4 > $O/plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4.csv $CMD4 > $O/ncu_list_c4.log 2>&1
echo "ncu list c4 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_prox_q -s 8 -c 1 -f -o $O/prof_c4 $CMD4 > $O/ncu_full_c4.log 2>&1
echo "ncu full c4 rc=$?"
CMDM="python bench.py --method 2 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --set full --clock-control none -k regex:k_tri_cluster -s 40 -c 1 -f -o $O/prof_tri $CMDM > $O/ncu_tri.log 2>&1
echo "ncu tri rc=$?"
