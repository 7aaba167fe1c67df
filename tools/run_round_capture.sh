#!/bin/sh
# One gpurun call: GPU tests, bench (every workload, both arms), launch lists and full ncu capture of the step kernels.
# usage (from the repo root on the GPU box): sh tools/run_round_capture.sh <tag>
TAG=${1:-r2}
O=${OUT_DIR:-gpurun_out}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > $O/${TAG}_gpu_tests.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench_driver_window.json 2>> $O/${TAG}_bench.err
for w in cfg2 cfg3 cfg5; do python bench.py --workload $w --steps 40 --warmup 5 > $O/${TAG}_bench_$w.json 2>> $O/${TAG}_bench.err; done
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_ref.json 2>> $O/${TAG}_bench.err
[ -x tools/ubench/host_shift ] && { tools/ubench/host_shift 4096 16 > $O/${TAG}_host_shift.txt 2>&1; tools/ubench/host_shift 4096 32 >> $O/${TAG}_host_shift.txt 2>&1; }
[ -x tools/ubench/lat ] && { tools/ubench/lat > $O/${TAG}_fp64_latency.txt 2>&1; tools/ubench/fp64tp >> $O/${TAG}_fp64_latency.txt 2>&1; }
# end-to-end step with host assembly: fixed splits and the handle's own balance (tools/time_e2e.py)
python tools/time_e2e.py 4096 > $O/${TAG}_e2e_host_assembly.txt 2>&1
# launch list of the step as the bench runs it (one graph replay per step: order + 16 x (physics, geometry, raster) + clear)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_graph.csv \
    python bench.py --steps 6 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launch_graph.log 2>&1
# kernel-level captures run the plain physics -> geometry -> raster sequence (MG_STEP_PARTS=1): under ncu every launch is
# serialised, and the 16 part launches of the overlapped step (each a latency-bound 16-CTA physics grid) would no longer
# add up to the step
MG_STEP_PARTS=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 12 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1
MG_STEP_PARTS=1 ncu --set full --clock-control none --import-source on -k regex:"mg_(physics|geometry|render)" -s 330 -c 3 -o $O/${TAG} \
    python bench.py --steps 12 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1
# physics section timers (clock64 build of the library)
if [ -f variants/libtimers.so ]; then
    MG_LIB_PATH=variants/libtimers.so MG_STEP_PARTS=1 python tools/section_timers.py MoveToCorner-Demo-LoRes4E-v0 4096 40 > $O/${TAG}_physics_sections.txt 2>&1
    MG_LIB_PATH=variants/libtimers.so MG_STEP_PARTS=1 python tools/section_timers.py FindDupe-TestLayout-LoRes4E-v0 16384 40 >> $O/${TAG}_physics_sections.txt 2>&1
    MG_LIB_PATH=variants/libtimers.so MG_STEP_PARTS=1 python tools/section_timers.py ClusterColour-TestAll-LoRes4E-v0 16384 40 >> $O/${TAG}_physics_sections.txt 2>&1
fi
cat $O/${TAG}_gpu_tests.log $O/${TAG}_bench.json
