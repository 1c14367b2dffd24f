#!/bin/bash
# round 2: per-agent cache-live flag + automatic 255-register cohorts: parity, then bench lines
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cached.py tests/test_gpu_free_running.py tests/test_gpu_full_size.py tests/test_gpu_ac.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r02c_pytest.log
cat gpurun_out/r02c_pytest.log
fmt='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4e"%d["value"], "%.4e"%d["e2e"]["value"], "frac %.3f"%d["roofline"]["frac"])'
for w in pendulum sweep cart_pole cart_double_pole_ac; do
python bench.py --workload $w --no-cpu --steps 5 --warmup 3 | python -c "$fmt" "$w"
done 2>&1 | tee gpurun_out/r02c_bench.txt
