#!/bin/bash
# Round 2, GPU call AY (1 GPU): ncu evidence of the final Method 2 step (launch list restricted to the step's kernels, --set full of the
# block dataflow solve); each ncu pass after the plain command exited 0
O=gpurun_out/r02_ay; mkdir -p $O
CMDM="python bench.py --method 2 --steps 1 --warmup 1 --no-cpu-baseline"
$CMDM > $O/plain_m2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_tri_sf|k_spmv|k_bi_|k_dot2|k_scalars|k_m2_|k_euler_elem|k_tri_level|k_finalize" -c 300 --csv --log-file $O/launches_m2.csv $CMDM > $O/ncu_list_m2.log 2>&1
echo "ncu list m2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_tri_sf_block -s 4 -c 2 -f -o $O/prof_tri $CMDM > $O/ncu_tri.log 2>&1
echo "ncu tri rc=$?"
