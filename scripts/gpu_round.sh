#!/bin/bash
# usage: gpu_round.sh <tag>  -- full GPU test suite, then the profile set (bench lines, ncu launch list, ncu full captures)
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest.log
tail -5 $O/pytest.log
bash scripts/gpu_profile.sh $TAG
for f in bench_c2 bench_c4 bench_ref; do echo "== $f"; tail -c 3000 $O/$f.json; tail -3 $O/$f.err; done
