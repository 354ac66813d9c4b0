#!/bin/bash
# Round 2, GPU call C (1 GPU): fixed tests; write-back probe (host time per queueing call, device time per copy)
O=gpurun_out/r02_c; mkdir -p $O
echo skip pytest
MMADMM_DEBUG_WB=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra-workloads > $O/bench_c2_snapkernel.json 2> $O/bench_c2_snapkernel.err; echo "rc=$?"; cat $O/bench_c2_snapkernel.err
MMADMM_DEBUG_WB=1 MMADMM_WB_SNAP=ce timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra-workloads > $O/bench_c2_snapce.json 2> $O/bench_c2_snapce.err; echo "rc=$?"; cat $O/bench_c2_snapce.err
MMADMM_DEBUG_WB=1 timeout 600 python bench.py --workload c3 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo "rc=$?"; cat $O/bench_c3.err
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra-workloads > $O/bench_c2_nodebug.json 2> $O/bench_c2_nodebug.err; echo "rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_c/bench_*.json')):
    try:
        d=json.load(open(f)); print(f, 'dev %.4f e2e %.4f strict %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['strict']['ms_per_step']), {k:round(v,4) for k,v in d['e2e']['phase_ms_per_step'].items()})
    except Exception as e: print(f, e)
PY
