for cfg in "16384 256 0" "16384 1024 0" "65536 256 0" "65536 64 0" "8192 256 0" "4096 256 0"; do
set -- $cfg
python bench.py --steps 2 --warmup 2 --no-cpu --agents $1 --memory $2 --cohort $3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$cfg', '%.3e'%d['value'], 'ms/pass %.2f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
done
