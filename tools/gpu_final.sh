#!/bin/bash
# What the driver runs at round end, on the final code: the whole GPU suite, smoke(), both bench arms, the training line.
O=gpurun_out; T=${1:-final}; mkdir -p $O
timeout 900 python -m pytest tests -x -q -m gpu > $O/${T}_pytest.log 2>&1; echo "gpu suite rc=$?"
tail -n 4 $O/${T}_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_ref.json 2> $O/${T}_ref.err; echo "reference arm rc=$?"
timeout 400 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --workload train > $O/${T}_train.json 2> $O/${T}_train.err; echo "train rc=$?"
python - $T <<'PY'
import json, sys
t = sys.argv[1]
b = json.load(open(f"gpurun_out/{t}_bench.json")); r = json.load(open(f"gpurun_out/{t}_ref.json")); tr = json.load(open(f"gpurun_out/{t}_train.json"))
print("value", b["value"], "e2e", b["e2e"]["value"], "ref", r["value"], "train ms", tr["ms_per_step"], "launches", b["gpu_launches"], "clocks", b["clocks"])
PY
