#!/bin/bash
# Round 2, GPU call T (1 GPU): mmadmm_energy_noalloc test alone in a fresh process, then the whole GPU suite (the new file runs
# last), then the evidence of the final library: bench lines, launch lists, ncu --set full of the prox / x-update kernels.
O=gpurun_out/r02_t; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 $O/pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -s --durations=12 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python bench.py --steps 20 --warmup 5 > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench (default line) rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?"
python bench.py --workload c4 --steps 10 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err; echo "bench c4 rc=$?"
python bench.py --workload c3 --steps 10 --warmup 5 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo "bench c3 rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra-workloads"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv $CMD > $O/ncu_list.log 2>&1
echo "ncu list c2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_prox_p -s 8 -c 1 -f -o $O/prof_c2 $CMD > $O/ncu_full.log 2>&1
echo "ncu full c2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_xupdate -s 12 -c 2 -f -o $O/prof_xup $CMD > $O/ncu_xup.log 2>&1
echo "ncu xup rc=$?"
CMD4="python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline"
$CMD4 > $O/plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c4.csv $CMD4 > $O/ncu_list_c4.log 2>&1
echo "ncu list c4 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_prox_q -s 8 -c 1 -f -o $O/prof_c4 $CMD4 > $O/ncu_full_c4.log 2>&1
echo "ncu full c4 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r02_t/bench_c2.json').read().split('\n') if l.startswith('{')][0])
w=d['workloads']
print('c2 %.4f e2e %.4f | c3 %.4f e2e %.4f | c4 %.4f e2e %.4f' % (d['ms_per_step'], d['e2e']['ms_per_step'], w['c3']['ms_per_step'], w['c3']['e2e']['ms_per_step'], w['c4']['ms_per_step'], w['c4']['e2e']['ms_per_step']))
print(d['phase_ms_per_step'], d['roofline'])
PY
