#!/bin/bash
# Round 2, GPU call V (1 GPU): block rows with short last group, device ILU, where-the-error-comes-from, parts timing
O=gpurun_out/r02_v; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_pcg.py tests/test_zz_gpu_method2.py -m gpu -q -s --durations=8 > $O/pytest_m2.log 2>&1; echo "pytest m2 rc=$?"; tail -14 $O/pytest_m2.log
timeout 600 python scripts/m2_parts.py 1024 > $O/m2_parts_nx1024.json 2> $O/m2_parts.err; echo "parts rc=$?"; cat $O/m2_parts_nx1024.json; tail -3 $O/m2_parts.err
for rb in 1 2; do
  timeout 900 python bench.py --method 2 --steps 5 --warmup 1 --no-cpu-baseline --opt m2_row_block=$rb > $O/bench_m2_rb$rb.json 2> $O/bench_m2_rb$rb.err; echo "m2 rb=$rb rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_v/bench_m2_rb*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
        print(f.split('/')[-1], '%.2f ms/step' % d['ms_per_step'], 'first step %.0f ms' % d['warmup_steps'][0]['ms'], d['config']['newton_iters'], d['config']['bicgstab_iters'], d['gpu_launches'], d['energy_last'])
    except Exception as e:
        print(f, 'ERR', e)
PY
