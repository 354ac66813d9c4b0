#!/bin/bash
# Round 2, GPU call A (1 GPU): full GPU test suite, default bench + reference arm, sanitizers on the small configs, ncu launch list.
O=gpurun_out/r02_a; mkdir -p $O
nproc > $O/nproc.txt; free -g > $O/mem.txt; nvidia-smi -L > $O/gpus.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q --durations=25 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -c 600 $O/bench.err
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/sanitize_small.py > $O/sanitizer_memcheck_1gpu.log 2>&1; echo "memcheck rc=$?"
timeout 1200 compute-sanitizer --tool racecheck --error-exitcode 9 python scripts/sanitize_small.py > $O/sanitizer_racecheck_1gpu.log 2>&1; echo "racecheck rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_c2.csv \
   python bench.py --steps 4 --warmup 3 --no-extra-workloads --no-cpu-baseline > $O/ncu_c2.log 2>&1; echo "ncu rc=$?"
