#!/bin/bash
# Round 2, GPU call R (1 GPU): step tail (speculative first solve + fused final reduction in k_xupdate): tests, then A/B
O=gpurun_out/r02_r; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for cfg in "step_tail=0" "step_tail=1" "step_tail=2" "step_tail=3" "step_tail=3 xup_variant=1" "step_tail=3 xup_variant=7" "step_tail=3 xup_variant=8"; do
  tag=$(echo $cfg | tr ' =' '__')
  opts=""; for o in $cfg; do opts="$opts --opt $o"; done
  timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline $opts > $O/bench_$tag.json 2> $O/bench_$tag.err; echo "$tag rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_r/bench_*.json')):
    try:
        d=json.loads([l for l in open(f).read().split('\n') if l.startswith('{')][0])
        w=d['workloads']
        print(f.split('/')[-1], 'c2 %.4f e2e %.4f xup %.4f fin %.4f | c3 %.4f | c4 %.4f xup %.4f' % (d['ms_per_step'], d['e2e'].get('ms_per_step',-1), d['phase_ms_per_step']['xupdate'], d['phase_ms_per_step'].get('residual',-1), w['c3']['ms_per_step'], w['c4']['ms_per_step'], w['c4']['phase_ms_per_step']['xupdate']))
    except Exception as e:
        print(f, 'ERR', e)
PY
