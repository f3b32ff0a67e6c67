#!/bin/bash
# GPU-box profiling pass for the POD path (run under gpurun): plain run first, then the ncu launch list,
# then one --set full capture of the dominant kernels. Outputs land in gpurun_out/.
TAG=${1:-r1}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-projection"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gram_kernel|lift_kernel|eig_jacobi|gram_reduce' -c 6 -f -o gpurun_out/prof_${TAG}_pod $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -n 2 gpurun_out/ncu_launch_$TAG.log gpurun_out/ncu_full_$TAG.log
