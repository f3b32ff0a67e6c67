#!/bin/bash
# Final-build profiling pass of round 2 (one GPU, short): launch list of the headline POD command and one full-set capture of
# gram_kernel and lift_kernel on one resident C3 chunk. Each ncu pass runs only after the same command exited 0 without ncu.
TAG=${1:-r2w}
O=gpurun_out
POD="python bench.py --steps 1 --warmup 1 --no-e2e --no-c2 --no-projection --no-checks --no-cpu-baseline"
$POD > $O/plain_pod_$TAG.log 2>&1 &&
timeout 170 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:ib200:: -c 3000 --csv --log-file $O/launches_pod_$TAG.csv $POD > $O/ncu_launch_pod_$TAG.log 2>&1
CH="$POD --cells 2500000"
$CH > $O/plain_chunk_$TAG.log 2>&1 &&
for K in gram_kernel lift_kernel; do
  timeout 100 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$K -c 1 --kill 1 -f -o /tmp/prof_${TAG}_pod_$K $CH > $O/ncu_full_pod_${K}_$TAG.log 2>&1
  ncu -i /tmp/prof_${TAG}_pod_$K.ncu-rep --page raw --csv > $O/raw_${TAG}_pod_$K.csv 2>/dev/null
done
ls -la $O | grep $TAG
