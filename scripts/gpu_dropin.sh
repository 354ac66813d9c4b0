#!/bin/bash
# usage: gpu_dropin.sh <tag> -- drop-in driver tests + an exploratory run of the Shoulder input (Monitor110)
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_dropin.py tests/test_abi.py -x -q -m gpu > $O/pytest_dropin.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest_dropin.log
W=/tmp/shoulder; mkdir -p $W/Experiments/InputFiles
python - <<PY
import json
t=json.load(open("tests/golden/ih_traces.json"))["Monitor110"]
s=dict(t["input"]); s["nSteps"]=19; s["DtTol"]=0
json.dump(s, open("$W/Experiments/InputFiles/Monitor110.json","w"))
print("ref", t["Ih"][:20])
PY
(cd $W && timeout 300 $GRAFT_REPO_ROOT/build/dropin/mesh.exe Monitor110 0 1 > run.log 2>&1; echo "shoulder rc=$?"; tail -3 run.log; cat Experiments/Results/Monitor110/Ih0.txt | awk -F, '{printf "%s ", $2}'; echo)
cp $W/run.log $O/shoulder.log 2>/dev/null
