#!/bin/bash
# Builds libscb200 variants with different onesweep tile sizes (HERE), and times one bench step
# per variant on the GPU box:   bash tools/sort_variants.sh build "a:-DSC_SORT_KPT_U32=4 -DSC_SORT_KPT_U64=16" ...
#                               gpurun -- 'bash tools/sort_variants.sh run'
cd "$(dirname "$0")/.."
D=street_crafter_b200
if [ "$1" = build ]; then
  shift; mkdir -p $D/_variants
  for spec in "$@"; do
    name=${spec%%:*}; defs=${spec#*:}
    for f in scan_sort projection; do
      extra=""; [ $f = projection ] && extra="-fmad=false"
      nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $extra $defs \
        -c $D/csrc/$f.cu -o $D/_variants/${f}_$name.o || exit 1
    done
    objs=$(ls $D/_obj/*.o | grep -v "scan_sort.o\|projection.o")
    nvcc -shared -o $D/_variants/libscb200_$name.so $objs $D/_variants/scan_sort_$name.o $D/_variants/projection_$name.o \
      -gencode arch=compute_100a,code=sm_100a && echo "built $name ($defs)"
  done
else
  mkdir -p gpurun_out
  for so in $D/_variants/libscb200_*.so; do
    echo "== $so"; SCB200_LIB=$PWD/$so python tools/sort_bench.py 2>&1 | grep "n=   626000\|REAL\|u64 n=\|u32 n="
    SCB200_LIB=$PWD/$so python tools/timeline.py 2 2>&1 | grep "steps:\|onesweep_pass_kernel<"
  done | tee gpurun_out/sort_variants.log
fi
