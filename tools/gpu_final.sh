#!/bin/bash
# What the driver runs at round end, on the final code: the whole GPU suite, smoke(), both bench arms.
O=gpurun_out; T=${1:-final}; mkdir -p $O
timeout 900 python -m pytest tests -x -q -m gpu > $O/${T}_pytest.log 2>&1; echo "gpu suite rc=$?"
tail -n 3 $O/${T}_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_ref.json 2> $O/${T}_ref.err; echo "reference arm rc=$?"
timeout 600 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
python - $T <<'PY'
import json, sys
t = sys.argv[1]
b = json.load(open(f"gpurun_out/{t}_bench.json")); r = json.load(open(f"gpurun_out/{t}_ref.json"))
print("value", b["value"], "graph", b["graph"]["value"], "e2e", b["e2e"]["value"], "ref", r["value"], "launches", b["gpu_launches"], "clocks", b["clocks"])
print("parity", {k: b["parity"][k] for k in ("max_abs", "psnr_db", "lists_bit_exact", "blend_bit_exact_on_same_gaussians")}, b["parity"]["oracle_vs_upstream_arithmetic"]["n_contrib_pixels_differ"])
print("T6", b["latency_T6"]["eager_ms_per_call"], b["latency_T6"]["graph_ms_per_call"], b["latency_T6"]["kernel_ms_per_call"])
print("train", b["train"]["ms_per_step"], b["train"].get("eager"), "stress", b["stress"]["value"], b["stress"]["ms_per_step"])
print("yard", {k: (v.get("ms_per_call"), v.get("render_loop_ms_per_call")) for k, v in b.get("gpu_yardsticks", {}).items() if isinstance(v, dict)}, b.get("gpu_yardsticks", {}).get("ours_render_batch_ms_per_call"))
print("cpu", b["cpu_baseline"])
for k, v in b["roofline"]["stages"].items():
    print(f"  {k:32s} {v['ms']:8.4f} ms  frac {v['frac']}")
PY
