#!/usr/bin/env bash
# Compile the REFERENCE's own CUDA kernels for sm_100a, from the sources where they lie
# under /root/reference, into oracle/_ref/libgsplat_ref.so (git-ignored, travels to the GPU
# box with the snapshot).  Flags = the reference's setup.py (gsplat/setup.py:76,85):
#   -O3 --use_fast_math --expt-relaxed-constexpr
# No reference source is copied; bindings.cu (needs torch/pybind11) is replaced by the
# C-ABI launcher oracle/ref_launch.cu.  TEST INFRASTRUCTURE ONLY.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${GS2D_REFERENCE_ROOT:-/root/reference}/gsplat/gsplat/cuda/csrc"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "build_ref: $REF not present (GPU box uses the prebuilt oracle/_ref)"; exit 0
fi
mkdir -p "$OUT"
FLAGS=(-O3 --use_fast_math --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a
       -Xcompiler -fPIC -I"$REF" -I"$REF/third_party/glm" -diag-suppress 550)
pids=()
for f in forward foward2d backward backward2d; do
  if [ ! -f "$OUT/$f.o" ] || [ "$REF/$f.cu" -nt "$OUT/$f.o" ]; then
    nvcc "${FLAGS[@]}" -c "$REF/$f.cu" -o "$OUT/$f.o" & pids+=($!)
  fi
done
nvcc "${FLAGS[@]}" -c "$HERE/ref_launch.cu" -o "$OUT/ref_launch.o" & pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
nvcc -shared -o "$OUT/libgsplat_ref.so" "$OUT"/forward.o "$OUT"/foward2d.o "$OUT"/backward.o "$OUT"/backward2d.o "$OUT"/ref_launch.o
echo "build_ref: wrote $OUT/libgsplat_ref.so"
