#!/bin/bash
O=gpurun_out/r02_ad; mkdir -p $O
for bs in 2 1; do timeout 300 python scripts/m2_sf_prof.py 1024 $bs > $O/sf_prof_bs$bs.log 2>&1; echo "bs=$bs rc=$?"; cat $O/sf_prof_bs$bs.log | tail -16; done
