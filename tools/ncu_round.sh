#!/bin/bash
# Profiling pass of one round (run on the GPU box through gpurun): launch lists + --set full captures.
# Every profiled command first exits 0 WITHOUT ncu; numbers printed under ncu are never bench values.
O=gpurun_out
SEQ="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras"
TRN="python bench.py --workload train --steps 2 --warmup 3"
OURS='regex:preprocess|scan_|scatter_entries|sort_tiles|sort_groups|blend|project|sample_decode|lbs_|pose_blend|skin|densify|chain_bwd|image_loss|camera'
$SEQ > $O/ncu_seq_plain.json 2> $O/ncu_seq_plain.err || exit 1
$TRN > $O/ncu_trn_plain.json 2> $O/ncu_trn_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 400 --csv \
    --log-file $O/launches_seq.csv $SEQ > $O/ncu_seq.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 600 --csv \
    --log-file $O/launches_trn.csv $TRN > $O/ncu_trn.log 2>&1
# full captures: one eager sequence step = 15 library launches (model set-up launches 1 more); skip set-up + 3 warm-up steps
ncu --set full --clock-control none --import-source on -k "$OURS" --launch-skip 46 -c 15 -f -o $O/seq_full $SEQ > $O/ncu_seq_full.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:^blend_bwd_kernel|^blend_kernel|sort_tiles|sort_groups|image_loss' \
    --launch-skip 40 -c 12 -f -o $O/trn_full $TRN > $O/ncu_trn_full.log 2>&1
ls -la $O | grep -E "full|launches"
