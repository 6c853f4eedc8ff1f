#!/usr/bin/env bash
# Compiles the REFERENCE's own perturbator sources, from where they lie under /root/reference,
# together with oracle/ref_harness.cu (which only calls them) into oracle/_ref/ref_harness.
# Nothing is copied into the repo; oracle/_ref/ is git-ignored but travels to the GPU box.
# Not built: planning_state.cu / main.cu (need ROS headers), the reference's CMake build.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${DOAPP_REFERENCE:-/root/reference}/perturbator"
OUT="$HERE/_ref"
[ -d "$REF" ] || { echo "reference tree not mounted at $REF" >&2; exit 0; }
mkdir -p "$OUT"
if [ "$OUT/ref_harness" -nt "$HERE/ref_harness.cu" ]; then exit 0; fi
NVCC="${NVCC:-$(command -v nvcc || echo /usr/local/cuda/bin/nvcc)}"
"$NVCC" -std=c++14 -O2 -gencode arch=compute_100a,code=sm_100a -rdc=true \
  -I "$REF/include" -I "$REF/nanoflann/include" \
  "$HERE/ref_harness.cu" "$REF/src/generate_trajectories.cu" "$REF/src/cpu_distance_grid.cu" \
  "$REF/src/common.cu" -o "$OUT/ref_harness" -lcurand -lpthread -Xcompiler -march=x86-64-v3
echo "built $OUT/ref_harness"
