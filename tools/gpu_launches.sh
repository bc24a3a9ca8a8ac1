#!/bin/bash
# ncu launch list (per-kernel durations) of a short bench run
# gpurun --timeout 900 -- 'bash tools/gpu_launches.sh tag'
TAG=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
