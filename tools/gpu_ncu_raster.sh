#!/bin/bash
# ncu --set full of one launch each of the compositing / SH / projection-backward / emit kernels
# gpurun --timeout 900 -- 'bash tools/gpu_ncu_raster.sh tag'
TAG=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on \
    -k regex:'rasterize_(fwd|bwd)|sh_bwd|sh_fwd|projection_bwd|projection_fwd|isect_emit' -s 42 -c 7 -o gpurun_out/${TAG}_raster -f \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_raster.log 2>&1
echo "ncu exit $?"
ls -la gpurun_out | tail
