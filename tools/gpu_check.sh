#!/bin/bash
# GPU evidence of a round, one ncu invocation per gpurun call.  Usage (from the repo root):
#   gpurun --timeout 1500 -- 'bash tools/gpu_check.sh <tag> 1'   GPU parity tests, the bench, the ncu launch list
#                                                                 of the same bench command
#   gpurun --timeout 1500 -- 'bash tools/gpu_check.sh <tag> 2'   one --set full capture of one launch of every
#                                                                 hot kernel
# then here:  python tools/ncu_summary.py <tag>   (writes profiles/<tag>_ncu_summary.txt)
TAG=${1:-r02}
PART=${2:-1}
mkdir -p gpurun_out
if [ "$PART" = "1" ]; then
  python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
  tail -3 gpurun_out/${TAG}_pytest.log
  python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
  echo "bench exit $?"; tail -c 300 gpurun_out/${TAG}_bench.json
  python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
  echo "reference arm exit $?"; tail -c 300 gpurun_out/${TAG}_bench_reference.json
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cfg4 > gpurun_out/${TAG}_plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv \
      --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cfg4 \
      > gpurun_out/${TAG}_ncu_launches.log 2>&1
  echo "ncu launches exit $?"
else
  # --set full of one launch of every hot kernel, from the EAGER sequential schedule (the same kernels the graph
  # replays, in a predictable launch order): per camera pass the filter matches 9 launches (projection fwd,
  # cooperative prepare, SH fwd, emit + column sort, row sort, the two compositing kernels, SH bwd, projection
  # bwd); skip 6 warm-up passes, capture one pass
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cfg4 --schedule sequential > gpurun_out/${TAG}_plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on \
      -k regex:'rasterize_|bin_prepare_coop|isect_emit_sort_tx|isect_sort_ty_final|sh_bwd|sh_fwd|projection_' -s 54 -c 9 \
      -o gpurun_out/${TAG}_prof -f \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cfg4 --schedule sequential > gpurun_out/${TAG}_ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
ls -la gpurun_out | tail -8
