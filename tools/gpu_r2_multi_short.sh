#!/bin/bash
# N-GPU lines with the final build: weak scaling of config 2, strong scaling of config 5 at its BASELINE size, reference arm
N=${1:-8}; tag=${2:-r2s}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:3}" > gpurun_out/${tag}_$2_n$N.json 2>> gpurun_out/${tag}_n$N.err; echo "$2 rc=$?"; }
run 29521 bench --steps 10 --warmup 3 --no-cpu-baseline
run 29522 bench_cfg5 --config 5 --global-batch 65536 --steps 3 --warmup 3 --no-cpu-baseline
run 29525 bench_cfg4 --config 4 --global-batch 4096 --steps 5 --warmup 3 --no-cpu-baseline
run 29526 bench_reference --impl reference --steps 3 --warmup 1
python - <<PY
import json
for n in ("bench","bench_cfg5","bench_cfg4","bench_reference"):
    try:
        d=json.load(open("gpurun_out/${tag}_%s_n$N.json" % n))
        print(n, "value %.0f" % d["value"], "e2e %.0f" % d["e2e"]["value"], d.get("scaling"), "gather_ms", d.get("gather_ms"), "comm", d.get("comm_bytes_per_step"), "ms/step %.2f" % d["ms_per_step"], d.get("status_counts"))
    except Exception as e:
        print(n, "ERR", e)
PY
tail -2 gpurun_out/${tag}_n$N.err
