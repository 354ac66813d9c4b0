#!/bin/bash
# Profiling build of the library for the timing experiments on the triangular solves (scripts/m2_prof_modes.py, m2_sf_prof.py):
# same sources with -DMMADMM_TRI_PROF (timing-only modes, per-chunk time stamps, debug exports).  Not a product build.
set -e
cd "$(dirname "$0")/.."
mkdir -p build/prof
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fopenmp -shared -DMMADMM_TRI_PROF \
  -o build/prof/libmmadmm_b200_prof.so mm-admm_b200/csrc/mmadmm.cu mm-admm_b200/csrc/partition.cu mm-admm_b200/csrc/pcg.cu -ldl -lgomp
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/prof/hop scripts/micro/hop.cu
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/prof/relay scripts/micro/relay.cu
echo "built build/prof/{libmmadmm_b200_prof.so,hop,relay}"
