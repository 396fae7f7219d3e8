#!/bin/bash
# Quick GPU check of a kernel change: the named test files, then the bench line (stages + 8f blocks).
O=gpurun_out; T=${1:-x}; shift
mkdir -p $O
timeout 400 python -m pytest "$@" -x -q -m gpu > $O/${T}_pytest.log 2>&1; echo "tests rc=$?" | tee -a $O/${T}_rc.txt
tail -n 8 $O/${T}_pytest.log
timeout 300 python bench.py --no-cpu-baseline > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?" | tee -a $O/${T}_rc.txt
tail -n 5 $O/${T}_bench.err
python - $T <<'PY'
import json, sys
b = json.load(open(f"gpurun_out/{sys.argv[1]}_bench.json"))
print(json.dumps({k: b.get(k) for k in ("value", "ms_per_step")}), {k: round(v["kernel_ms"], 4) for k, v in b.get("producer_layout", {}).items() if isinstance(v, dict)})
print({k: v["ms"] for k, v in b["roofline"]["stages"].items()})
PY
