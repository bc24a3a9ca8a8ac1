#!/bin/bash
# gpurun --timeout 900 -- 'bash tools/pipe_variants_run.sh'   : cfg2 step under the schedules of pipeline.ViewPipeline
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/pipe_test.log 2>&1; echo "test=$?"; tail -3 gpurun_out/pipe_test.log
run() {
  tag=$1; shift
  timeout 300 python bench.py --no-cpu-baseline --no-cfg4 "$@" > gpurun_out/pipe_$tag.json 2> gpurun_out/pipe_$tag.err || { echo "$tag FAILED"; tail -5 gpurun_out/pipe_$tag.err; return; }
  python - "$tag" <<'PY'
import json, sys
d = json.load(open("gpurun_out/pipe_%s.json" % sys.argv[1]))
print(sys.argv[1], "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), "step_frac", round(d["roofline_step"]["frac"], 4),
      {k[3:]: round(v["ms_per_step"], 3) for k, v in d["stages"].items() if isinstance(v, dict)})
PY
}
run seq --schedule sequential
run a1 --schedule pipeline --ahead 1
run a3 --schedule pipeline --ahead 3
run a1_notail --schedule pipeline --ahead 1 --no-pipe-tail
run a3_nosh --schedule pipeline --ahead 3 --no-pipe-sh
