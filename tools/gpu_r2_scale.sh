#!/bin/bash
# weak-scaling line of config 2 at N GPUs with the final build
N=${1:-2}; tag=${2:-r2x}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_n$N.json 2> gpurun_out/${tag}_n$N.err; echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench_n$N.json')); print('N=$N value %.0f e2e %.0f gather_ms %s ms/step %.2f' % (d['value'], d['e2e']['value'], d.get('gather_ms'), d['ms_per_step']))"
