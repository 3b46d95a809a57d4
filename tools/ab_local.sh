#!/bin/bash
# A/B with local-memory counters: tools/ab_local.sh "<variants>"  (ncu metrics only on B=1024)
for v in $1; do
  if [ $v = cur ]; then unset NMPC_LIB; else export NMPC_LIB=$PWD/iterative_learning_nmpc_b200/libnmpc_b200_$v.so; fi
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --metrics smsp__sass_inst_executed_op_local_ld.sum,smsp__sass_inst_executed_op_local_st.sum,l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct,smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:nmpc_qp_fused -s 3 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | grep -E "local|inst_executed.sum|gpu__time" | sed "s/^/$v /"
done
