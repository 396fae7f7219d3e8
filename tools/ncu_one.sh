#!/bin/bash
# One --set full capture of one kernel (regex $1) of the sequence bench at $3 frames (default 30), tag $2.
# The plain run goes first (exit code checked), the profiled run repeats the same command.
O=gpurun_out; K=$1; T=${2:-k}; FR=${3:-30}; SKIP=${4:-3}
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --frames $FR ${BENCH_ARGS}"
$CMD > $O/${T}_plain.json 2> $O/${T}_plain.err &&
ncu --set full --clock-control none --import-source on -k "regex:$K" -s $SKIP -c 1 -f -o $O/${T}_full $CMD > $O/${T}_ncu.log 2>&1
echo "ncu rc=$?"; tail -n 3 $O/${T}_ncu.log; ls -la $O/${T}_full.ncu-rep
