#!/bin/bash
# Round 2, GPU call AV (1 GPU): 3D Method 2 preconditioner per mode; host set-up phases of the first Method 2 step; Method 2 bench line
O=gpurun_out/r02_av; mkdir -p $O
timeout 600 python scripts/m2_parts3d.py 40 > $O/m2_parts3d_nx40.json 2> $O/parts3d.err; echo "parts3d rc=$?"; cat $O/m2_parts3d_nx40.json; tail -2 $O/parts3d.err
MMADMM_DEBUG_SETUP=1 timeout 600 python bench.py --method 2 --steps 5 --warmup 1 --no-cpu-baseline > $O/bench_m2.json 2> $O/bench_m2.err; echo "bench m2 rc=$?"; grep "mmadmm setup" $O/bench_m2.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_av/bench_m2.json').read().split('\n') if l.startswith('{')][0])
print('m2', d['ms_per_step'], d['warmup_steps'])
PY
