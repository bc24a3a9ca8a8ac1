#!/bin/bash
# gpurun -- 'bash tools/graph_variants_run.sh' : graphed two-stream step under compositing-occupancy limits
for w in 0 14 12 10 8; do for a in 1 3; do
  echo "== RASTER_WARPS=$w ahead=$a"; RASTER_WARPS=$w timeout 200 python tools/graph_probe.py 2000000 $a 2>&1 | grep "replay [45]\|Error" ; done; done
echo "== prio 0, warps 12"; MEM_PRIO=0 RASTER_WARPS=12 timeout 200 python tools/graph_probe.py 2000000 1 2>&1 | grep "replay [45]\|Error"
