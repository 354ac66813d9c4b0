#!/bin/bash
# Round 2, GPU call Y (1 GPU): dataflow (sync-free) triangular solve
O=gpurun_out/r02_y; mkdir -p $O
timeout 900 python -m pytest tests/test_zz_gpu_method2.py -m gpu -q -s -x -k "block_row" > $O/pytest_blk.log 2>&1; echo "pytest blk rc=$?"; tail -8 $O/pytest_blk.log
timeout 1200 python -m pytest tests/test_gpu_pcg.py tests/test_zz_gpu_method2.py -m gpu -q -s --durations=5 > $O/pytest_m2.log 2>&1; echo "pytest m2 rc=$?"; tail -8 $O/pytest_m2.log
timeout 600 python scripts/m2_parts.py 1024 > $O/m2_parts_nx1024.json 2> $O/m2_parts.err; echo "parts rc=$?"; cat $O/m2_parts_nx1024.json; tail -3 $O/m2_parts.err
for cfg in "m2_tri_mode=2 m2_row_block=1" "m2_tri_mode=2 m2_row_block=2"; do
  tag=$(echo $cfg | tr ' =' '__'); opts=""; for o in $cfg; do opts="$opts --opt $o"; done
  timeout 900 python bench.py --method 2 --steps 5 --warmup 1 --no-cpu-baseline $opts > $O/bench_$tag.json 2> $O/bench_$tag.err; echo "$tag rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_y/bench_m2*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
        print(f.split('/')[-1], '%.2f ms/step' % d['ms_per_step'], 'first step %.0f ms' % d['warmup_steps'][0]['ms'], d['config']['newton_iters'], d['config']['bicgstab_iters'], d['gpu_launches'], d['energy_last'])
    except Exception as e:
        print(f, 'ERR', e)
PY
