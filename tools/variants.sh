#!/bin/bash
# Generic tuning-variant builder:  FILES="sh" bash tools/variants.sh build "name:-DFOO=1" "other:-DBAR"
# rebuilds the listed csrc files with the extra defines and links street_crafter_b200/_variants/libscb200_<name>.so
# (run a tool against one with SCB200_LIB=<path>).
cd "$(dirname "$0")/.."
D=street_crafter_b200
[ "$1" = build ] || { echo "usage: FILES='a b' $0 build name:defs ..."; exit 1; }
shift; mkdir -p $D/_variants
for spec in "$@"; do
  name=${spec%%:*}; defs=${spec#*:}
  excl="XXXX"; objs_new=""
  for f in $FILES; do
    extra=""
    [ $f = projection ] && extra="-fmad=false"
    [ $f = rasterize ] && extra="--use_fast_math"
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $extra $defs \
      -Xptxas -v -c $D/csrc/$f.cu -o $D/_variants/${f}_$name.o 2> $D/_variants/${f}_$name.ptxas || { cat $D/_variants/${f}_$name.ptxas; exit 1; }
    excl="$excl\|/$f.o"; objs_new="$objs_new $D/_variants/${f}_$name.o"
  done
  objs=$(ls $D/_obj/*.o | grep -v "$excl")
  nvcc -shared -o $D/_variants/libscb200_$name.so $objs $objs_new -gencode arch=compute_100a,code=sm_100a && echo "built $name ($defs)"
done
