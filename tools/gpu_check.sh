#!/bin/bash
# One gpurun call: GPU parity tests, the bench, the ncu launch list of the same bench command and
# one --set full capture of one launch of every hot kernel.  Usage (from the repo root):
#   gpurun --timeout 1500 -- 'bash tools/gpu_check.sh [tag]'
TAG=${1:-r01}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -c 400 gpurun_out/${TAG}_bench.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
# per camera pass the filter matches 13 launches (projection fwd, 4+2 onesweep, emit, SH fwd, the two
# compositing kernels, SH bwd, projection bwd): skip 6 warm-up passes, capture one pass
ncu --set full --clock-control none --import-source on \
    -k regex:'rasterize_|onesweep|sh_bwd|sh_fwd|projection_|isect_emit' -s 78 -c 13 -o gpurun_out/${TAG}_prof -f \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
# the loss kernels run in the e2e leg only: capture them from the stand-alone loss tool
ncu --set full --clock-control none --import-source on -k regex:photometric -s 6 -c 2 \
    -o gpurun_out/${TAG}_loss -f python tools/loss_bench.py > gpurun_out/${TAG}_ncu_loss.log 2>&1
echo "ncu loss exit $?"
# the densification-cycle / compaction kernels from their stand-alone tools
python tools/densify_bench.py > gpurun_out/${TAG}_densify.log 2>&1; python tools/compact_bench.py >> gpurun_out/${TAG}_densify.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'densify_|compact_rows' -s 3 -c 3 \
    -o gpurun_out/${TAG}_densify -f python tools/densify_bench.py > gpurun_out/${TAG}_ncu_densify.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'compact_rows' -s 1 -c 1 \
    -o gpurun_out/${TAG}_compact -f python tools/compact_bench.py > gpurun_out/${TAG}_ncu_compact.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'assemble_' -s 6 -c 2 \
    -o gpurun_out/${TAG}_assemble -f python tools/assemble_bench.py > gpurun_out/${TAG}_ncu_assemble.log 2>&1
echo "ncu densify exit $?"; cat gpurun_out/${TAG}_densify.log | tail -3
ls -la gpurun_out | tail -12
