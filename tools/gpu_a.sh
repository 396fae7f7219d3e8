#!/bin/bash
# round-2 call A: new config-size parity tests, the whole GPU suite, the default bench, the FFMA2 microbenchmark
O=gpurun_out; T=${1:-a}; mkdir -p $O
nvidia-smi -L > $O/${T}_gpu.txt
./tools/ubench/ffma2 > $O/${T}_ffma2.txt 2>&1; cat $O/${T}_ffma2.txt
timeout 900 python -m pytest tests/test_config_parity_gpu.py -x -q -m gpu --durations=5 > $O/${T}_parity.log 2>&1; echo "parity rc=$?"; tail -n 12 $O/${T}_parity.log
timeout 900 python -m pytest tests -x -q -m gpu --deselect tests/test_config_parity_gpu.py > $O/${T}_pytest.log 2>&1; echo "gpu suite rc=$?"; tail -n 4 $O/${T}_pytest.log
timeout 400 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"; tail -c 600 $O/${T}_bench.json
