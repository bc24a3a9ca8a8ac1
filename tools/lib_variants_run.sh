#!/bin/bash
# gpurun -- 'bash tools/lib_variants_run.sh'  : the bench's per-stage table (eager sequential schedule) and the graph
# step time for the default library and every street_crafter_b200/_variants/libscb200_*.so (built HERE by
# tools/raster_variants.sh / tools/sort_variants.sh)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {
  tag=$1
  timeout 300 python bench.py --no-cpu-baseline --no-cfg4 > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err || { echo "$tag FAILED"; tail -3 gpurun_out/var_$tag.err; return; }
  python - "$tag" <<'PY'
import json, sys
d = json.load(open("gpurun_out/var_%s.json" % sys.argv[1]))
print(sys.argv[1], "graph ms/step", round(d["ms_per_step"], 4), "eager", round(d["config"]["ms_per_step_eager_sequential"], 4),
      {k[3:]: round(v["ms_per_step"], 4) for k, v in d["stages"].items() if isinstance(v, dict)})
PY
}
run default
shopt -s nullglob
for so in street_crafter_b200/_variants/libscb200_*.so; do
  name=$(basename $so .so); name=${name#libscb200_}
  SCB200_LIB=$PWD/$so run $name
done
