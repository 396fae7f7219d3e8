#!/bin/bash
# multi-GPU check ($1 = N): 2-GPU peer-gather test, then both bench arms under torchrun as the driver launches them
N=${1:-2}; T=${2:-n$N}; O=gpurun_out; mkdir -p $O
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_peer_gather_gpu.py -x -q -m gpu 2>&1 | tail -3
NCCL_DEBUG=INFO timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N ${BENCH_ARGS} > $O/${T}_bench.out 2> $O/${T}_bench.err; echo "bench rc=$?"
tail -n 1 $O/${T}_bench.out > $O/${T}_bench.json; grep -c "NCCL INFO" $O/${T}_bench.err; grep -m2 "nranks\|NVLS" $O/${T}_bench.err | cut -c1-200; wc -l $O/${T}_bench.out
grep -v 'NCCL INFO' $O/${T}_bench.err | tail -n 5 | cut -c1-300
python - $T <<'PY'
import json, sys
b = json.load(open(f"gpurun_out/{sys.argv[1]}_bench.json"))
print("value", round(b["value"], 1), "ms", round(b["ms_per_step"], 3), b["scaling"], "graph", b.get("graph"), "gather_check", b.get("gather_check"))
print("weak", b.get("weak")); print("e2e", b.get("e2e", {}).get("value"))
t = b.get("train", {}); print("train", t.get("value"), t.get("ms_per_step"), t.get("config", {}).get("jobs_this_rank"))
s = b.get("stress", {}); print("stress", s.get("value"), s.get("ms_per_step"), s.get("overflow"))
print("latency", b.get("latency_T6"))
PY
