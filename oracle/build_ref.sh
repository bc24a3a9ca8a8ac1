#!/bin/bash
# Builds oracle/_ref/libsimpleknn_ref.so from the reference's simple-knn sources IN PLACE
# (nothing is copied into this repo).  Runs only where /root/reference exists (the build
# container); the .so then travels to the GPU box with gpurun (oracle/_ref is git-ignored,
# not gpurun-ignored).  gsplat / diff_point_rasterization are absent from /root/reference and
# cannot be built.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference/submodules/simple-knn
[ -f "$REF/simple_knn.cu" ] || { echo "reference sources not present: skipping oracle/_ref"; exit 0; }
mkdir -p "$HERE/_ref"
OUT="$HERE/_ref/libsimpleknn_ref.so"
if [ "$OUT" -nt "$REF/simple_knn.cu" ] && [ "$OUT" -nt "$HERE/ref_knn_wrapper.cu" ]; then exit 0; fi
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC -w -I"$REF" \
      -shared -o "$OUT" "$REF/simple_knn.cu" "$HERE/ref_knn_wrapper.cu"
echo "built $OUT"
