#!/bin/bash
# ncu --set full capture of the fused QP kernel (one launch after warm-up) for the bench workload.
tag=${1:-r1}
cmd="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$cmd > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:nmpc_qp_fused -s 3 -c 1 -o gpurun_out/${tag}_fused $cmd > gpurun_out/${tag}_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/${tag}_ncu.log
