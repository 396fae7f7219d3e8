#!/bin/bash
# --set full capture of kernels matching $1 in the training step (2 launches of the last step), tag $2
O=gpurun_out; K=$1; T=${2:-tr}; SKIP=${3:-8}
CMD="python bench.py --workload train --steps 1 --warmup 3"
$CMD > $O/${T}_plain.json 2> $O/${T}_plain.err &&
ncu --set full --clock-control none --import-source on -k "regex:$K" -s $SKIP -c 2 -f -o $O/${T}_full $CMD > $O/${T}_ncu.log 2>&1
echo "ncu rc=$?"; tail -n 2 $O/${T}_ncu.log | cut -c1-200
