#!/bin/bash
# Final GPU pass of round 1 (run through gpurun): parity tests, bench lines, launch lists, one full capture.
# Every profiled command first exits 0 WITHOUT ncu; numbers printed under ncu are never bench values.
O=gpurun_out; P=${1:-n}
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/${P}_gpu.txt
timeout 240 python -m pytest tests/test_next_rows_gpu.py -q -m gpu > $O/${P}_pytest_new.log 2>&1; echo "new tests rc=$?" | tee -a $O/${P}_rc.txt
timeout 600 python -m pytest tests -x -q -m gpu > $O/${P}_pytest_gpu.log 2>&1; echo "gpu suite rc=$?" | tee -a $O/${P}_rc.txt
tail -3 $O/${P}_pytest_new.log $O/${P}_pytest_gpu.log
timeout 300 python bench.py > $O/${P}_bench.json 2> $O/${P}_bench.err; echo "bench rc=$?" | tee -a $O/${P}_rc.txt
timeout 200 python bench.py --workload train > $O/${P}_train.json 2> $O/${P}_train.err; echo "train rc=$?" | tee -a $O/${P}_rc.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/${P}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${P}_rc.txt
SEQ="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
TRN="python bench.py --workload train --steps 2 --warmup 3"
OURS='regex:preprocess|scan_|scatter_entries|sort_tiles|blend|project|sample_decode|lbs_|pose_blend|skin|densify|chain_bwd|image_loss|camera|frames_to_u8'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 400 --csv \
    --log-file $O/${P}_launches_seq.csv $SEQ > $O/${P}_ncu_seq.log 2>&1; echo "ncu seq rc=$?" | tee -a $O/${P}_rc.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 400 --csv \
    --log-file $O/${P}_launches_trn.csv $TRN > $O/${P}_ncu_trn.log 2>&1; echo "ncu trn rc=$?" | tee -a $O/${P}_rc.txt
timeout 400 ncu --set full --clock-control none --import-source on -k "$OURS" --launch-skip 61 -c 15 -f -o $O/${P}_seq_full $SEQ > $O/${P}_ncu_seq_full.log 2>&1; echo "ncu full rc=$?" | tee -a $O/${P}_rc.txt
cat $O/${P}_rc.txt; head -c 1500 $O/${P}_bench.json
