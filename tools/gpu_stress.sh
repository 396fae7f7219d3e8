#!/bin/bash
# Stress sizes (BASELINE.json configs[4]: 500k Gaussians, 2048^2) on one GPU: full-size property tests, the bench line
# profiles/r01_stress_*.json came from, and its ncu launch list.  Run through gpurun.
O=gpurun_out; T=${1:-stress}; mkdir -p $O
timeout 500 python -m pytest tests/test_full_size_gpu.py -q -m gpu > $O/${T}_pytest.log 2>&1; echo "full-size tests rc=$?"
B="python bench.py --frames 16 --gaussians 500000 --size 2048 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 400 $B > $O/${T}_stress.json 2> $O/${T}_stress.err; echo "stress rc=$?"
OURS='regex:preprocess|scan_|scatter_entries|sort_tiles|blend|project|sample_decode|lbs_|pose_blend|skin|densify|camera'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 400 --csv --log-file $O/${T}_launches_stress.csv $B > $O/${T}_ncu.log 2>&1; echo "ncu rc=$?"
head -c 600 $O/${T}_stress.json
