#!/bin/sh
# One gpurun call: GPU tests, bench (both arms), launch list and full ncu capture of the step kernels.
# usage (from the repo root on the GPU box): sh tools/run_round_capture.sh <tag>
TAG=${1:-r1c}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/${TAG}_gpu_tests.log
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
# kernel-level captures run the plain physics -> geometry -> raster sequence (MG_STEP_PARTS=1): under ncu every launch is
# serialised, and the 16 part launches of the overlapped step (each a latency-bound 16-CTA physics grid) would no longer
# add up to the step; ${TAG}_launches_parts.csv keeps the list of the overlapped form as launched by the bench
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_parts.csv \
    python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launch_parts.log 2>&1
MG_STEP_PARTS=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 12 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launch.log 2>&1
MG_STEP_PARTS=1 ncu --set full --clock-control none --import-source on -k regex:"mg_(physics|geometry|render)" -s 15 -c 3 -o gpurun_out/${TAG} \
    python bench.py --steps 12 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
cat gpurun_out/${TAG}_gpu_tests.log gpurun_out/${TAG}_bench.json
