#!/bin/bash
# The reference's own driver (main.cpp, unchanged; build/dropin/mesh.exe) on the C2 input: SquareGrid 1024^2, MEx4, 10 steps.
TAG=$1; O=$GRAFT_REPO_ROOT/gpurun_out/$TAG; mkdir -p $O
W=/tmp/drv; mkdir -p $W/Experiments/InputFiles
for nx in 256 1024; do
cat > $W/Experiments/InputFiles/Square$nx.json <<JSON
{"TestType": "SquareGrid", "Dim": 2, "MonType": 4, "Method": 0, "CompMesh": false, "BoundaryType": 1, "GradUse": false,
 "nSteps": 10, "AdmmIter": 10, "DtTol": 0, "dt": 0.025, "tau": 0.5, "rho": 50, "w": 3.53553390593,
 "nx": $nx, "ny": $nx, "xa": 0, "xb": 1, "ya": 0, "yb": 1}
JSON
( cd $W && T0=$(date +%s.%N) && $GRAFT_REPO_ROOT/build/dropin/mesh.exe Square$nx 0 1 > $O/driver_$nx.log 2> $O/driver_$nx.err; echo "nx=$nx rc=$?" )
grep -E "Took|Number of|Time spent" $O/driver_$nx.log; 
head -12 $W/Experiments/Results/Square$nx/Ih0.txt | tr '\n' ' '; echo
done
