#!/bin/bash
# final-build lines for configs 3-5 and the ncu launch list of the default bench (ncu last, after the plain run exited 0)
tag=${1:-run}
timeout 900 python bench.py --config 3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg3.json 2> gpurun_out/${tag}_bench.err; echo "cfg3 rc=$?"; cut -c1-200 gpurun_out/${tag}_bench_cfg3.json
timeout 900 python bench.py --config 4 --batch 4096 --steps 5 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg4.json 2>> gpurun_out/${tag}_bench.err; echo "cfg4 rc=$?"; cut -c1-200 gpurun_out/${tag}_bench_cfg4.json
timeout 900 python bench.py --config 5 --batch 16384 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg5.json 2>> gpurun_out/${tag}_bench.err; echo "cfg5 rc=$?"; cut -c1-200 gpurun_out/${tag}_bench_cfg5.json
tail -3 gpurun_out/${tag}_bench.err
cmd="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$cmd > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $cmd > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu rc=$?"
