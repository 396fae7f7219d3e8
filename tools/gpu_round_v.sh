#!/bin/bash
# Medium-bucket sort instantiation: binning parity tests, then c2 / c4 / stress bench lines.
O=gpurun_out; T=${1:-v}; mkdir -p $O
timeout 400 python -m pytest tests/test_splat_gpu.py tests/test_full_size_gpu.py tests/test_splat_bwd_gpu.py -x -q -m gpu > $O/${T}_pytest.log 2>&1; echo "tests rc=$?"
tail -n 6 $O/${T}_pytest.log
timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --workload train > $O/${T}_train.json 2> $O/${T}_train.err; echo "train rc=$?"
timeout 300 python bench.py --frames 16 --gaussians 500000 --size 2048 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > $O/${T}_stress.json 2> $O/${T}_stress.err; echo "stress rc=$?"
python - $T <<'PY'
import json, sys
for n in ("bench", "train", "stress"):
    b = json.load(open(f"gpurun_out/{sys.argv[1]}_{n}.json"))
    st = b["roofline"]["stages"] if "roofline" in b and "stages" in b["roofline"] else b.get("stages", {})
    print(n, round(b["value"], 1), round(b["ms_per_step"], 3), {k: v["ms"] for k, v in st.items() if "bin" in k or "blend" in k})
PY
