#!/bin/bash
# Builds variants of libscb200.so with different rasterize.cu knobs (HERE, no GPU needed):
#   bash tools/raster_variants.sh build "name1:-DSC_FWD_MIN_CTAS=5" "name2:-DSC_BWD_MIN_CTAS=5" ...
# and times them on the GPU box:  gpurun -- 'bash tools/raster_variants.sh run'
cd "$(dirname "$0")/.."
D=street_crafter_b200
if [ "$1" = build ]; then
  shift; mkdir -p $D/_variants
  for spec in "$@"; do
    name=${spec%%:*}; defs=${spec#*:}
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --use_fast_math $defs \
      -Xptxas -v -c $D/csrc/rasterize.cu -o $D/_variants/rast_$name.o 2> $D/_variants/rast_$name.log || { tail $D/_variants/rast_$name.log; exit 1; }
    objs=$(ls $D/_obj/*.o | grep -v rasterize.o)
    nvcc -shared -o $D/_variants/libscb200_$name.so $objs $D/_variants/rast_$name.o -gencode arch=compute_100a,code=sm_100a
    echo "$name: $(grep -A2 'rasterize_fwd_kernelILi4ELb0ELb0' $D/_variants/rast_$name.log | grep -o 'Used [0-9]* registers') / bwd $(grep -A2 'rasterize_bwd_kernelILi4' $D/_variants/rast_$name.log | grep -o 'Used [0-9]* registers')"
  done
else
  mkdir -p gpurun_out
  for so in $D/_variants/libscb200_*.so; do
    echo "== $so"; SCB200_LIB=$PWD/$so python tools/raster_sweep.py 2>&1 | grep -v "^yaw" 
  done | tee gpurun_out/raster_variants.log
fi
