#!/bin/bash
# Full-size property tests (configs[1] and configs[4] sizes) + a stress-size bench line (500k Gaussians, 2048^2).
O=gpurun_out; mkdir -p $O
timeout 500 python -m pytest tests/test_full_size_gpu.py -q -m gpu > $O/t_pytest.log 2>&1; echo "full-size tests rc=$?" | tee -a $O/t_rc.txt
tail -n 25 $O/t_pytest.log
timeout 400 python bench.py --frames 16 --gaussians 500000 --size 2048 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > $O/t_stress.json 2> $O/t_stress.err; echo "stress rc=$?" | tee -a $O/t_rc.txt
tail -n 5 $O/t_stress.err; head -c 1200 $O/t_stress.json
