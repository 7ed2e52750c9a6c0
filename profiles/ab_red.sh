timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
for v in 0 1 0 1; do
  GTK_WGRAD_RED=$v timeout 300 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/abr_$v.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/abr_$v.json')); print('red=$v graph', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])"
done
