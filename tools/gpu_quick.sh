#!/bin/bash
# quick GPU iteration: parity tests + bench (no ncu).  gpurun --timeout 900 -- 'bash tools/gpu_quick.sh tag'
TAG=${1:-q}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -15 gpurun_out/${TAG}_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
print("value",d["value"],"ms/step",d["ms_per_step"],"e2e",d["e2e"]["value"],"kernel_ms",d["roofline_step"]["kernel_ms_per_step"])
for k,v in d["stages"].items(): print(f"  {k:28s} {v['ms_per_step']:.3f} ms/step  calls {v['calls_per_step']}")
print(d["clocks"], d["roofline"]["frac"])
PY
