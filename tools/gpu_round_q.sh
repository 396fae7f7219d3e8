#!/bin/bash
# ncu captures of project_tm_kernel (v2), project_kernel and scatter_entries at bench size, with source counters.
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:project_tm_kernel -c 1 -f -o $O/q_tm_full $B > $O/q_ncu_tm.log 2>&1; echo "ncu tm rc=$?" | tee -a $O/q_rc.txt
timeout 200 ncu --set full --clock-control none --import-source on -k 'regex:scatter_entries|^project_kernel|project_kernel' --launch-skip 8 -c 2 -f -o $O/q_sc_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $O/q_ncu_sc.log 2>&1; echo "ncu sc rc=$?" | tee -a $O/q_rc.txt
