#!/bin/bash
# A/B of library builds on one box: tools/ab_probe.sh "<variant> ..." "<batch> ..."   (variant "cur" = the product library)
for v in $1; do
  if [ $v = cur ]; then unset NMPC_LIB; else export NMPC_LIB=$PWD/iterative_learning_nmpc_b200/libnmpc_b200_$v.so; fi
  for B in $2; do
    python bench.py --batch $B --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', $B, 'fused ms', round(d['roofline']['launch_ms'],3), 'value', round(d['value']), 'e2e', round(d['e2e']['value']))"
  done
done
