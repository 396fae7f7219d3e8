#!/bin/bash
# 2-GPU check of the final code: peer-gather parity test, then the bench lines the driver's scaling run launches.
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_peer_gather_gpu.py -x -q -m gpu > $O/n2_pytest.log 2>&1; echo "peer tests rc=$?" | tee -a $O/n2_rc.txt
tail -n 5 $O/n2_pytest.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/n2_bench.json 2> $O/n2_bench.err; echo "bench n2 rc=$?" | tee -a $O/n2_rc.txt
tail -n 5 $O/n2_bench.err; head -c 600 $O/n2_bench.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/n2_ref.json 2> $O/n2_ref.err; echo "ref n2 rc=$?" | tee -a $O/n2_rc.txt
head -c 600 $O/n2_ref.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --workload train --gpus 2 > $O/n2_train.json 2> $O/n2_train.err; echo "train n2 rc=$?" | tee -a $O/n2_rc.txt
head -c 400 $O/n2_train.json
