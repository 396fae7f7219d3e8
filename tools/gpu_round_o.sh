#!/bin/bash
# Second GPU pass: token-major projection (8f-2) tests + the bench line with the sink / producer-layout blocks.
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_next_rows_gpu.py tests/test_stages_gpu.py -q -m gpu > $O/o_pytest_new.log 2>&1; echo "new tests rc=$?" | tee -a $O/o_rc.txt
tail -n 15 $O/o_pytest_new.log
timeout 300 python bench.py > $O/o_bench.json 2> $O/o_bench.err; echo "bench rc=$?" | tee -a $O/o_rc.txt
tail -n 5 $O/o_bench.err
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:project_tm_kernel -c 1 -f -o $O/o_tm_full $B > $O/o_ncu_tm.log 2>&1; echo "ncu tm rc=$?" | tee -a $O/o_rc.txt
timeout 200 ncu --set full --clock-control none --import-source on -k regex:frames_to_u8 --launch-skip 60 -c 1 -f -o $O/o_sink_full $B > $O/o_ncu_sink.log 2>&1; echo "ncu sink rc=$?" | tee -a $O/o_rc.txt
python - <<'PY'
import json
b = json.load(open("gpurun_out/o_bench.json"))
print(json.dumps({k: b.get(k) for k in ("value", "e2e", "sink", "producer_layout")}, indent=1))
PY
