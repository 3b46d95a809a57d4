#!/bin/bash
# quick A/B line: config 2 at one full wave, at the bench batch and at four waves
mkdir -p gpurun_out
for b in 592 1024 2368; do
  python bench.py --batch $b --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('${1:-run} B $b', 'fused ms %.3f' % d['roofline'].get('launch_ms', -1), 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'], d.get('status_counts'))
" | tee -a gpurun_out/quick_ab.txt
done
