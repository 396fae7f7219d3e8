#!/bin/bash
# Third GPU pass: cp.async token-major projection + scatter_entries latency changes.
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_next_rows_gpu.py tests/test_splat_gpu.py tests/test_renderer_gpu.py -x -q -m gpu > $O/p_pytest.log 2>&1; echo "tests rc=$?" | tee -a $O/p_rc.txt
tail -n 15 $O/p_pytest.log
timeout 300 python bench.py > $O/p_bench.json 2> $O/p_bench.err; echo "bench rc=$?" | tee -a $O/p_rc.txt
tail -n 5 $O/p_bench.err
timeout 200 python bench.py --workload train > $O/p_train.json 2> $O/p_train.err; echo "train rc=$?" | tee -a $O/p_rc.txt
python - <<'PY'
import json
b = json.load(open("gpurun_out/p_bench.json"))
print(json.dumps({k: b.get(k) for k in ("value", "ms_per_step", "producer_layout")}, indent=1))
print({k: v["ms"] for k, v in b["roofline"]["stages"].items()})
t = json.load(open("gpurun_out/p_train.json")); print(t["value"], t["ms_per_step"])
PY
