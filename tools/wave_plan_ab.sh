#!/bin/bash
# A/B of the wave-by-wave launch plan of the fused QP kernel (NMPC_WAVE_PLAN=1, default) against one launch per chunk
mkdir -p gpurun_out
run() {  # label, env, args
  NMPC_WAVE_PLAN=$2 python bench.py ${@:3} --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('$1 plan=$2', 'ms_per_step %.3f' % d['ms_per_step'], 'fused ms %.3f' % d['roofline'].get('launch_ms', -1), 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'], d.get('status_counts', ''), 'launches', d.get('gpu_launches'))
" | tee -a gpurun_out/wave_plan_ab.txt
}
for plan in 0 1; do
  for b in 740 1024 2368 4096; do run "cfg2 B=$b" $plan --batch $b --steps 6 --warmup 3; done
  run "cfg3 B=1024" $plan --config 3 --batch 1024 --steps 3 --warmup 3
  run "cfg5 B=8192" $plan --config 5 --batch 8192 --steps 3 --warmup 3
done
