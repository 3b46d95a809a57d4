#!/bin/bash
N=8; tag=r2w
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:3}" > gpurun_out/${tag}_$2_n$N.json 2>> gpurun_out/${tag}_n$N.err; echo "$2 rc=$?"; }
run 29521 bench --steps 10 --warmup 3 --no-cpu-baseline
run 29522 bench_cfg5 --config 5 --global-batch 65536 --steps 3 --warmup 3 --no-cpu-baseline
python - <<PY
import json
for n in ("bench","bench_cfg5"):
    d=json.load(open("gpurun_out/${tag}_%s_n$N.json" % n))
    print(n, "value %.0f" % d["value"], "e2e %.0f" % d["e2e"]["value"], d.get("scaling"), "gather_ms", d.get("gather_ms"), "ms/step %.2f" % d["ms_per_step"], "e2e ms %.2f" % d["e2e"]["ms_per_step"])
PY
tail -2 gpurun_out/${tag}_n$N.err
