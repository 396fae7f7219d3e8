#!/bin/bash
# bin-stage iteration: the bit-exact list tests, then the stage table of the 120-frame sequence (and optionally the stress sizes)
O=gpurun_out; T=${1:-bin}; mkdir -p $O
AMAV_SKIP_SLOW=${SKIP_SLOW:-1} timeout 900 python -m pytest tests/test_splat_gpu.py tests/test_config_parity_gpu.py tests/test_full_size_gpu.py -x -q -m gpu > $O/${T}_pytest.log 2>&1; echo "tests rc=$?"; tail -n 3 $O/${T}_pytest.log
timeout 300 python bench.py --no-e2e --no-cpu-baseline --no-extras > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
python - $T <<'PY'
import json, sys
b = json.load(open(f"gpurun_out/{sys.argv[1]}_bench.json"))
print("value", b["value"], "ms", b["ms_per_step"], "clocks", b["clocks"]["sm_mhz"])
for k, v in b["roofline"]["stages"].items():
    print(f"  {k:32s} {v['ms']:8.4f} ms")
PY
