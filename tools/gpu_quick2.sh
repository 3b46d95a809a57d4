#!/bin/bash
# quick check of a kernel change: QP parity tests, then the default bench line
tag=${1:-q}
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "qp_iterates or without_inequalities or rti_solve or conditional" > gpurun_out/${tag}_pytest_qp.log 2>&1; echo "pytest-qp rc=$?"; tail -15 gpurun_out/${tag}_pytest_qp.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
python - <<'PY'
import json,sys
try:
    d=json.load(open("gpurun_out/'$tag'_bench.json".replace("'","")))
except Exception as e:
    print("no json", e); sys.exit(0)
print("value", d["value"], "e2e", d["e2e"]["value"], "fused launch ms", d["roofline"]["launch_ms"], "status", d.get("status_counts"))
PY
