#!/bin/bash
# Round 2, GPU call U (1 GPU): block-row triangular solves ("row_block"), device ILU(0) factor / refactor, sliding-test tolerance
O=gpurun_out/r02_u; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_pcg.py tests/test_zz_gpu_method2.py tests/test_zz_no_alloc.py -m gpu -q -x --durations=8 > $O/pytest_m2.log 2>&1; echo "pytest m2 rc=$?"; tail -12 $O/pytest_m2.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "sliding" > $O/pytest_sliding.log 2>&1; echo "pytest sliding rc=$?"; tail -3 $O/pytest_sliding.log
for rb in 1 2 4; do
  timeout 900 python bench.py --method 2 --steps 5 --warmup 1 --no-cpu-baseline --opt m2_row_block=$rb > $O/bench_m2_rb$rb.json 2> $O/bench_m2_rb$rb.err; echo "m2 rb=$rb rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_u/bench_m2_rb*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
        print(f.split('/')[-1], '%.2f ms/step' % d['ms_per_step'], 'first step %.0f ms' % d['warmup_steps'][0]['ms'], d['config']['newton_iters'], d['config']['bicgstab_iters'], d['gpu_launches'], d['energy_last'])
    except Exception as e:
        print(f, 'ERR', e)
PY
