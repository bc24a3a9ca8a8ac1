#!/bin/bash
# runs a tool against every built variant:  bash tools/run_variants.sh <glob-part> <python tool> [args]
cd "$(dirname "$0")/.."
pat=$1; shift
for so in street_crafter_b200/_variants/libscb200_${pat}*.so; do
  echo "== $so"; SCB200_LIB=$PWD/$so python "$@" 2>&1 | tail -6
done
