#!/bin/bash
# generic round-2 GPU check: whole GPU suite (config-size parity included), then the default bench line
O=gpurun_out; T=${1:-b}; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu --durations=8 > $O/${T}_pytest.log 2>&1; echo "gpu suite rc=$?"; tail -n 15 $O/${T}_pytest.log
timeout 400 python bench.py ${BENCH_ARGS} > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"; tail -n 3 $O/${T}_bench.err
python - $T <<'PY'
import json, sys
b = json.load(open(f"gpurun_out/{sys.argv[1]}_bench.json"))
print("value", round(b["value"], 1), "ms", round(b["ms_per_step"], 3), "e2e", round(b.get("e2e", {}).get("value", 0), 1), "launches", b["gpu_launches"], b["clocks"])
for k, v in b["roofline"]["stages"].items():
    print(f"  {k:32s} {v['ms']:8.4f} ms  frac {v['frac']}")
PY
