#!/bin/bash
# Launch list of the stress sizes (500k Gaussians, 2048^2, 16 frames).
O=gpurun_out; mkdir -p $O
B="python bench.py --frames 16 --gaussians 500000 --size 2048 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
OURS='regex:preprocess|scan_|scatter_entries|sort_tiles|blend|project|sample_decode|lbs_|pose_blend|skin|densify|camera'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 400 --csv --log-file $O/u_launches_stress.csv $B > $O/u_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import sys
sys.path.insert(0, "tools")
import ncu_summarise as N
print(N.launch_summary("gpurun_out/u_launches_stress.csv", "stress"))
PY
