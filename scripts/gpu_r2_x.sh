#!/bin/bash
# Round 2, GPU call X (1 GPU): where a level of the cluster triangular solve spends its time (profiling build, timing-only modes)
O=gpurun_out/r02_x; mkdir -p $O
for bs in 1 2; do for m in 0 1 2 3 4; do
  MMADMM_TRI_PROF_MODE=$m timeout 300 python scripts/m2_prof_modes.py 1024 $bs >> $O/modes.jsonl 2>> $O/modes.err; echo "bs=$bs mode=$m rc=$?"
done; done
cat $O/modes.jsonl; tail -3 $O/modes.err
