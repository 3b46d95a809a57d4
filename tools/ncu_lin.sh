#!/bin/bash
# ncu --set full capture of the linearisation kernels (one launch each after warm-up) for the bench workload.
tag=${1:-r1}
cmd="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$cmd > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"nmpc_linearize_sr|nmpc_model_kernel" -s 6 -c 2 -o gpurun_out/${tag}_lin $cmd > gpurun_out/${tag}_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/${tag}_ncu.log
