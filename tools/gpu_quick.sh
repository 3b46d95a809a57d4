#!/bin/bash
# quick GPU pass: parity tests, config-2 bench line, optional phase timing (needs libnmpc_b200_timing.so)
tag=${1:-q}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench.json"))
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "fused ms", round(d["roofline"]["launch_ms"],3), "frac", round(d["roofline"]["frac"],3), d["status_counts"])
PY
if [ -f iterative_learning_nmpc_b200/libnmpc_b200_timing.so ]; then
  NMPC_LIB=$PWD/iterative_learning_nmpc_b200/libnmpc_b200_timing.so python tests/gpu_fused_timing.py 148 > gpurun_out/${tag}_timing_148.txt 2>&1
  tail -21 gpurun_out/${tag}_timing_148.txt
fi
