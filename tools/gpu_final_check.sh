#!/bin/bash
# last check of the tree ($TAG): the whole GPU suite, smoke(), the default bench line and the reference arm
TAG=${TAG:-r02final}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_${TAG}.log
tail -3 gpurun_out/pytest_gpu_${TAG}.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${TAG}.log 2>&1; echo "smoke exit $?"
python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref exit $?"
cat gpurun_out/bench_${TAG}.json
