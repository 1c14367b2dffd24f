#!/bin/bash
# one --set full capture of the run kernel per workload in $NCU_WORKLOADS, summarised on the box ($TAG)
TAG=${TAG:-r02x}
mkdir -p gpurun_out
for w in ${NCU_WORKLOADS:-pendulum}; do
  python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_run -s 6 -c 1 -o gpurun_out/prof_${TAG}_$w \
      python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full_$w.log 2>&1
  { echo "# ncu --set full --clock-control none --import-source on -k regex:k_run -s 6 -c 1 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu"
    echo "# (tools/gpu_ncu_one.sh, TAG=$TAG; one launch = 100 agent-steps per agent of the launch's cohort; times under ncu are not bench values)"
    python tools/ncu_summary.py gpurun_out/prof_${TAG}_$w.ncu-rep
    echo; echo "# per source line (tools/ncu_by_line.py)"
    python tools/ncu_by_line.py gpurun_out/prof_${TAG}_$w.ncu-rep grl_b200/csrc/libgrlcu.so 30 2>&1
  } > gpurun_out/ncu_${TAG}_$w.txt
  [ "$KEEP_REP" = "$w" ] || rm -f gpurun_out/prof_${TAG}_$w.ncu-rep
done
