#!/bin/bash
# A/B of nmpc_linearize_sr_kernel builds on one box: tools/ab_lin_sr.sh "<variant> ..."   ("cur" = product library)
for v in $1; do
  if [ $v = cur ]; then unset NMPC_LIB; else export NMPC_LIB=$PWD/iterative_learning_nmpc_b200/libnmpc_b200_$v.so; fi
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['per_kind']
print('$v', 'linearize ms', round(k['linearize']['launch_ms'],4), 'fused ms', round(k['qp_fused']['launch_ms'],3), 'value', round(d['value']), 'e2e', round(d['e2e']['value']))"
done
