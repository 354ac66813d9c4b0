#!/bin/bash
# Round 2, GPU call AJ (1 GPU): evidence of the final library -- whole GPU suite, bench lines (default line with c3/c4 records,
# reference arm, Method 2 line), ncu launch list of a Method 2 step and ncu --set full of the dataflow triangular-solve kernels.
O=gpurun_out/r02_aj; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q -s --durations=12 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench (default line) rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
python bench.py --method 2 --steps 5 --warmup 1 > $O/bench_m2.json 2> $O/bench_m2.err; echo "bench method 2 rc=$?"
python bench.py --method 2 --steps 5 --warmup 1 --no-cpu-baseline --opt m2_tri_mode=1 --opt m2_row_block=1 --opt m2_tri_layout=0 > $O/bench_m2_round1_kernels.json 2> $O/bench_m2_round1_kernels.err; echo "bench method 2 (cluster kernels, row per thread) rc=$?"
CMDM="python bench.py --method 2 --steps 1 --warmup 1 --no-cpu-baseline"
$CMDM > $O/plain_m2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_m2.csv $CMDM > $O/ncu_list_m2.log 2>&1
echo "ncu list m2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_tri_sf_block -s 4 -c 2 -f -o $O/prof_tri $CMDM > $O/ncu_tri.log 2>&1
echo "ncu tri rc=$?"
python - <<'PY'
import json
for f in ('bench_c2','bench_m2','bench_m2_round1_kernels','bench_ref'):
    try:
        d=json.loads([l for l in open('gpurun_out/r02_aj/%s.json'%f).read().split('\n') if l.startswith('{')][0])
        print(f, d['value'], d['unit'], d['ms_per_step'] if 'ms_per_step' in d else '', d.get('e2e',{}).get('value'))
    except Exception as e: print(f,'ERR',e)
PY
