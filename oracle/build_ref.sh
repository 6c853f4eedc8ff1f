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
NVCC="${NVCC:-$(command -v nvcc || echo /usr/local/cuda/bin/nvcc)}"
# Stage-2 pin (ref_fk_harness.cu): the reference's generate_trajectories.cu compiled INTO the
# harness TU so its static __device__ helpers (:303-387) are reachable, with the one line :394
# (`return 0; //Quick hack`) dropped so num_collisions (:391-466) runs.  The edited copy is a
# build product under the git-ignored _ref/ directory.
if [ ! "$OUT/ref_fk_harness" -nt "$HERE/ref_fk_harness.cu" ]; then
  sed '394{/Quick hack/d}' "$REF/src/generate_trajectories.cu" > "$OUT/gt_nohack.cu"
  if cmp -s "$OUT/gt_nohack.cu" "$REF/src/generate_trajectories.cu"; then
    echo "build_ref.sh: line 394 of generate_trajectories.cu is not the expected stub" >&2; exit 1
  fi
  "$NVCC" -std=c++14 -O2 -gencode arch=compute_100a,code=sm_100a -rdc=true \
    -I "$REF/include" -I "$REF/nanoflann/include" -DGT_SOURCE="\"$OUT/gt_nohack.cu\"" \
    "$HERE/ref_fk_harness.cu" "$REF/src/cpu_distance_grid.cu" "$REF/src/common.cu" \
    -o "$OUT/ref_fk_harness" -lcurand -lpthread -Xcompiler -march=x86-64-v3
  echo "built $OUT/ref_fk_harness"
fi
if [ "$OUT/ref_harness" -nt "$HERE/ref_harness.cu" ]; then exit 0; fi
"$NVCC" -std=c++14 -O2 -gencode arch=compute_100a,code=sm_100a -rdc=true \
  -I "$REF/include" -I "$REF/nanoflann/include" \
  "$HERE/ref_harness.cu" "$REF/src/generate_trajectories.cu" "$REF/src/cpu_distance_grid.cu" \
  "$REF/src/common.cu" -o "$OUT/ref_harness" -lcurand -lpthread -Xcompiler -march=x86-64-v3
echo "built $OUT/ref_harness"
