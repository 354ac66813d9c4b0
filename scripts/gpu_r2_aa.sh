#!/bin/bash
O=gpurun_out/r02_aa; mkdir -p $O
for g in 0 1 2 37; do MMADMM_SF_GRID=$g timeout 300 python scripts/dbg_sf.py >> $O/dbg.log 2>&1; echo "grid $g rc=$?"; done
cat $O/dbg.log
