#!/bin/bash
# round-2 evidence pass: GPU tests, bench lines of every config, launch list and ncu full capture of the fused kernel
tag=${1:-r2k}
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print(\"smoke ok\")" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${tag}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/${tag}_bench.json
timeout 600 python bench.py --impl reference > gpurun_out/${tag}_bench_reference.json 2>> gpurun_out/${tag}_bench.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/${tag}_bench_reference.json
timeout 900 python bench.py --config 3 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg3.json 2>> gpurun_out/${tag}_bench.err; echo "cfg3 rc=$?"
timeout 900 python bench.py --config 3 --qp-warm-start 2 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg3_hot.json 2>> gpurun_out/${tag}_bench.err; echo "cfg3 hot rc=$?"
timeout 900 python bench.py --config 4 --batch 4096 --steps 5 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg4.json 2>> gpurun_out/${tag}_bench.err; echo "cfg4 rc=$?"
timeout 900 python bench.py --config 5 --batch 16384 --steps 3 --no-cpu-baseline > gpurun_out/${tag}_bench_cfg5.json 2>> gpurun_out/${tag}_bench.err; echo "cfg5 rc=$?"
timeout 900 python bench.py --config 1 > gpurun_out/${tag}_bench_cfg1.json 2>> gpurun_out/${tag}_bench.err; echo "cfg1 rc=$?"
python - <<PY
import json
for n in ("bench","bench_cfg3","bench_cfg3_hot","bench_cfg4","bench_cfg5","bench_cfg1"):
    try:
        d=json.load(open("gpurun_out/${tag}_%s.json" % n))
        print(n, "value %.0f e2e %.0f" % (d["value"], d["e2e"]["value"]), "status", d.get("status_counts"), "sqp hist", d.get("sqp_iter_histogram"), d.get("tick_ms"))
    except Exception as e:
        print(n, "ERR", e)
PY
cmd="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $cmd > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:nmpc_qp_fused -s 3 -c 1 -o gpurun_out/${tag}_fused $cmd > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu full rc=$?"
