#!/bin/bash
# evidence pass ($TAG): parity, smoke, bench lines of every workload, reference arm, ncu launch list of the default
# bench command and one --set full capture of the run kernel per workload, summarised on the box
TAG=${TAG:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_${TAG}.log
tail -3 gpurun_out/pytest_gpu_${TAG}.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${TAG}.log 2>&1; echo "smoke exit $?"
python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref exit $?"
for w in cart_pole cart_double_pole_ac sweep cart_double_pole pendulum_ac cart_pole_ac; do
  python bench.py --workload $w --no-cpu > gpurun_out/bench_${w}_${TAG}.json 2> gpurun_out/bench_${w}_${TAG}.err; echo "$w exit $?"
done
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${TAG}.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
for w in ${NCU_WORKLOADS:-pendulum cart_pole cart_double_pole_ac sweep}; do
  python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_run -s 6 -c 1 -o gpurun_out/prof_${TAG}_$w \
      python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full_$w.log 2>&1
  # summarise on the box (gpurun_out/ is capped at 64 MiB and one report is ~27 MB): counters + per-source-line attribution
  { echo "# ncu --set full --clock-control none --import-source on -k regex:k_run -s 6 -c 1 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu"
    echo "# (tools/gpu_evidence.sh, TAG=$TAG; one launch = 100 agent-steps per agent of the launch's cohort; times under ncu are not bench values)"
    python tools/ncu_summary.py gpurun_out/prof_${TAG}_$w.ncu-rep
    echo; echo "# per source line (tools/ncu_by_line.py)"
    python tools/ncu_by_line.py gpurun_out/prof_${TAG}_$w.ncu-rep grl_b200/csrc/libgrlcu.so 25 2>&1
  } > gpurun_out/ncu_${TAG}_$w.txt
  [ "$KEEP_REP" = "$w" ] || rm -f gpurun_out/prof_${TAG}_$w.ncu-rep
done
ls -la gpurun_out | head -60
