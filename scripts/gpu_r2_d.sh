#!/bin/bash
O=gpurun_out/r02_d; mkdir -p $O
run() { name=$1; shift; env "$@" timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra-workloads $EXTRA > $O/$name.json 2> $O/$name.err; echo "== $name rc=$?"; grep "stream synchronise\|copy duration" $O/$name.err; }
EXTRA="" run base MMADMM_DEBUG_WB=1
EXTRA="--no-e2e-phases" run nophase MMADMM_DEBUG_WB=1
EXTRA="" run chunks16 MMADMM_DEBUG_WB=1 MMADMM_WB_CHUNKS=16
EXTRA="--no-e2e-phases" run chunks16_nophase MMADMM_DEBUG_WB=1 MMADMM_WB_CHUNKS=16
EXTRA="--workload c3" run c3_base MMADMM_DEBUG_WB=1
EXTRA="--workload c3 --no-e2e-phases" run c3_chunks16_nophase MMADMM_DEBUG_WB=1 MMADMM_WB_CHUNKS=16
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_d/*.json')):
    try:
        d=json.load(open(f)); print(f, 'dev %.4f e2e %.4f strict %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['strict']['ms_per_step']))
    except Exception as e: print(f, e)
PY
