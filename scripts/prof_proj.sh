#!/bin/bash
# GPU-box profiling pass for the projection path (200k-cell mesh, N=50): launch list + one --set full capture.
TAG=${1:-r1}
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --projection c4small"
$CMD > gpurun_out/plainp_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:ib200:: -c 1500 --csv --log-file gpurun_out/launches_proj_$TAG.csv $CMD > gpurun_out/ncu_launchp_$TAG.log 2>&1
$CMD > gpurun_out/plainp2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'conv_contract|conv_face_prep|div_phi_kernel|laplacian_kernel|grad_vec_kernel|face_flux|bnd_curl|grad_scalar' -c 14 -f -o gpurun_out/prof_${TAG}_proj $CMD > gpurun_out/ncu_fullp_$TAG.log 2>&1
tail -n 3 gpurun_out/ncu_launchp_$TAG.log gpurun_out/ncu_fullp_$TAG.log
