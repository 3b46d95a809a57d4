#!/bin/bash
# One GPU-box pass: parity tests, bench (both arms), extra workload lines.  Outputs under gpurun_out/<tag>_*.
tag=${1:-run}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/${tag}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json | cut -c1-600
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2>> gpurun_out/${tag}_bench.err; echo "ref rc=$?"
timeout 600 python bench.py --config 3 --batch 4096 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg3.json 2>> gpurun_out/${tag}_bench.err; echo "cfg3 rc=$?"
cut -c1-400 gpurun_out/${tag}_bench_cfg3.json
timeout 600 python bench.py --batch 16384 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_b16k.json 2>> gpurun_out/${tag}_bench.err; echo "b16k rc=$?"
cut -c1-400 gpurun_out/${tag}_bench_b16k.json
tail -5 gpurun_out/${tag}_bench.err
