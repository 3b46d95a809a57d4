#!/bin/bash
# round-2 first pass: GPU parity tests, bench lines of configs 2/3/1, ncu launch list
tag=${1:-r2a}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; cat gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
timeout 900 python bench.py --config 3 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg3.json 2> gpurun_out/${tag}_bench_cfg3.err; echo "cfg3 rc=$?"; cat gpurun_out/${tag}_bench_cfg3.json; tail -3 gpurun_out/${tag}_bench_cfg3.err
timeout 900 python bench.py --config 1 > gpurun_out/${tag}_bench_cfg1.json 2> gpurun_out/${tag}_bench_cfg1.err; echo "cfg1 rc=$?"; cat gpurun_out/${tag}_bench_cfg1.json; tail -3 gpurun_out/${tag}_bench_cfg1.err
cmd="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $cmd > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu rc=$?"
