#!/bin/bash
# launch time of the fused QP kernel against the number of co-resident CTAs per SM (148 SMs): is a wave bound by the
# chain latency of one CTA or by what the co-resident CTAs share?
mkdir -p gpurun_out
for b in 148 296 444 592 740 888 1024; do
  python bench.py --batch $b --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('B', $b, 'ms_per_step %.3f' % d['ms_per_step'], 'fused launch ms %.3f' % d['roofline'].get('launch_ms', -1), 'solves/s %.0f' % d['value'])
" | tee -a gpurun_out/residency_sweep.txt
done
