#!/bin/bash
# N-GPU pass (gpurun --gpus N): NCCL gather check + weak-scaling bench lines, both arms, launched as the driver does
n=${1:-2}; tag=${2:-multi}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR tools/gpu_multi_check.py > gpurun_out/${tag}_check_n$n.log 2>&1; echo "check rc=$?"; grep "world=" gpurun_out/${tag}_check_n$n.log
timeout 600 $TR bench.py --impl reference --gpus $n --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference_n$n.json 2> gpurun_out/${tag}_n$n.err; echo "ref rc=$?"
timeout 600 $TR bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n$n.json 2>> gpurun_out/${tag}_n$n.err; echo "bench rc=$?"
python - <<PY
import json
for f in ("gpurun_out/${tag}_bench_reference_n$n.json", "gpurun_out/${tag}_bench_n$n.json"):
    d = json.loads([l for l in open(f) if l.startswith("{")][-1])
    print(f, d.get("impl", "native"), "n_gpus", d["n_gpus"], "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d.get("clocks"))
PY
tail -3 gpurun_out/${tag}_n$n.err
