for cfg in "65536 65536 0" "65536 65536 32768" "65536 65536 16384" "65536 65536 8192" "131072 65536 0" "32768 65536 0" "16384 65536 0"; do
set -- $cfg
python bench.py --steps 3 --warmup 2 --no-cpu --agents $1 --memory $2 --cohort $3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('$cfg', '%.3e'%d['value'], 'e2e %.3e'%d['e2e']['value'], 'ms/pass %.2f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'], 'V %.1f'%d['roofline']['V_mean'])"
done
