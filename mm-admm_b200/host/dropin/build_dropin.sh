#!/bin/bash
# Builds the reference's own driver (main.cpp, UNCHANGED, compiled from where it lies under $REF) against the
# drop-in headers of this directory and links it with libmmadmm_b200.so -> build/dropin/mesh.exe.
#
# main.cpp includes "./src/Mesh.h", "./lib/json/...", "Experiments/TestMonitors/MEx*.h" relative to its own
# directory, so a staging tree of symlinks is assembled: main.cpp and the monitor plug-ins point into the reference,
# src/, lib/json and Eigen/ point at this directory.  No reference source is copied into the repository.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(cd "$HERE/../../.." && pwd)"
REF="${REF:-/root/reference}"
OUT="${OUT:-$ROOT/build/dropin}"
if [ ! -f "$REF/main.cpp" ]; then echo "reference not present at $REF: keeping prebuilt $OUT/mesh.exe"; exit 0; fi
# The staging tree lives in a temporary directory outside the repository and is removed after the compile.
mkdir -p "$OUT"; rm -rf "$OUT/stage"
STAGE="$(mktemp -d /tmp/mmadmm_dropin_stage.XXXXXX)"
trap 'rm -rf "$STAGE"' EXIT
mkdir -p "$STAGE/Experiments/TestMonitors" "$STAGE/lib"
ln -s "$REF/main.cpp" "$STAGE/main.cpp"
for f in "$REF"/Experiments/TestMonitors/*.h; do ln -s "$f" "$STAGE/Experiments/TestMonitors/$(basename "$f")"; done
ln -s "$HERE/src" "$STAGE/src"
ln -s "$HERE/lib/json" "$STAGE/lib/json"
g++ -O2 -std=c++17 -w -I"$HERE" -I"$ROOT/include" "$STAGE/main.cpp" -o "$OUT/mesh.exe" \
    -L"$ROOT/mm-admm_b200" -lmmadmm_b200 -Wl,-rpath,'$ORIGIN/../../mm-admm_b200'
echo "built $OUT/mesh.exe"
