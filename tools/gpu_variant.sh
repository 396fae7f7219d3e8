#!/bin/bash
# A/B of one source file's compile-time switch on the box: tools/gpu_variant.sh <file.cu> "<-D flags>" <tag> [extra nvcc flags]
# rebuilds that object with the flags, relinks, and prints the stage table of the 120-frame sequence.
F=$1; D=$2; T=$3; X=$4; P=audio_motion_avatar_b200
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I include -I $P/csrc $X $D -c $P/csrc/$F -o $P/build/${F%.cu}.o &&
nvcc -shared -o $P/libamav_b200.so $P/build/*.o -lcudart || { echo "build failed"; exit 1; }
timeout 300 python bench.py --no-e2e --no-cpu-baseline ${BENCH_EXTRA:---no-extras} > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python - $T "$D" <<'PY'
import json, sys
b = json.load(open(f"gpurun_out/{sys.argv[1]}_bench.json"))
s = b["roofline"]["stages"]
for k in ("train", "stress"):
    if k in b: print(" ", k, {q: b[k].get(q) for q in ("value", "ms_per_step")}, {q[5:]: round(v["ms"], 3) for q, v in (b[k].get("stages_one_call") or {}).items()})
print(sys.argv[2], "value", round(b["value"], 1), "ms", round(b["ms_per_step"], 3), " ".join(f"{k[5:]}={v['ms']:.3f}" for k, v in s.items()))
PY
