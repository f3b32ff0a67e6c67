#!/bin/bash
# Round-2 GPU-box profiling pass (run under gpurun, one GPU). Every ncu pass runs only after the same command has exited 0
# without ncu. Outputs land in gpurun_out/; scripts/ncu_summary.py turns them into profiles/*.md and the traffic JSONs.
#   1. launch list of the headline command (C3 on one GPU, row-streamed) -> the kernels' shares of the step
#   2. --set full of gram_kernel / lift_kernel / the eigensolve kernels on one resident C3 chunk (2.5M cells x 1000 vector snapshots)
#   3. launch list of the C4 projection + one --set full capture PER KERNEL (-k regex, -c 2) so that the dominant kernels
#      (conv_contract, conv_face_prep, div_phi, face_flux, grad_scalar) are all captured
TAG=${1:-r2}
O=gpurun_out
POD="python bench.py --steps 1 --warmup 1 --no-e2e --no-c2 --no-projection --no-checks --no-cpu-baseline"
$POD > $O/plain_pod_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:ib200:: -c 3000 --csv --log-file $O/launches_pod_$TAG.csv $POD > $O/ncu_launch_pod_$TAG.log 2>&1
CH="$POD --cells 2500000"
$CH > $O/plain_chunk_$TAG.log 2>&1 &&
# gpurun brings back at most 64 MiB: the reports are turned into raw-page CSVs on the box and removed
for K in gram_kernel lift_kernel eig_jacobi gram_reduce; do
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$K -c 1 --kill 1 -f -o /tmp/prof_${TAG}_pod_$K $CH > $O/ncu_full_pod_${K}_$TAG.log 2>&1
  ncu -i /tmp/prof_${TAG}_pod_$K.ncu-rep --page raw --csv > $O/raw_${TAG}_pod_$K.csv 2>/dev/null
done
ncu -i /tmp/prof_${TAG}_pod_gram_kernel.ncu-rep --page source --csv > $O/src_${TAG}_pod_gram_kernel.csv 2>/dev/null
export ITHACA_BENCH_MESH_CACHE=/tmp/c4mesh
PRJ="python scripts/proj_bench.py c4"
$PRJ > $O/plain_proj_$TAG.json 2> $O/plain_proj_$TAG.err &&
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:ib200:: -c 3000 --csv --log-file $O/launches_proj_$TAG.csv $PRJ > $O/ncu_launch_proj_$TAG.log 2>&1
for K in conv_contract_kernel conv_face_prep_kernel div_phi_kernel face_flux_kernel grad_scalar_kernel laplacian_kernel grad_vec_kernel surface_integrate_kernel; do
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$K -c 2 --kill 1 -f -o /tmp/prof_${TAG}_proj_$K $PRJ > $O/ncu_full_proj_${K}_$TAG.log 2>&1
  ncu -i /tmp/prof_${TAG}_proj_$K.ncu-rep --page raw --csv > $O/raw_${TAG}_proj_$K.csv 2>/dev/null
done
ncu -i /tmp/prof_${TAG}_proj_conv_contract_kernel.ncu-rep --page source --csv > $O/src_${TAG}_proj_conv_contract_kernel.csv 2>/dev/null
ncu -i /tmp/prof_${TAG}_proj_laplacian_kernel.ncu-rep --page source --csv > $O/src_${TAG}_proj_laplacian_kernel.csv 2>/dev/null
du -sh $O
ls -la $O | grep $TAG | head -40
