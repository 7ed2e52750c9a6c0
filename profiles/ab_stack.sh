for v in 0 1 0 1; do
  GTK_WN_STACK=$v python bench.py --no-cpu-baseline --steps 20 > gpurun_out/ab_$v.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/ab_$v.json')); print('stack=$v graph', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'inf', round(d['inference']['ms_per_call'],3), 'launches', d['gpu_launches'])"
done
for v in 0 1; do
  GTK_WN_STACK=$v python bench.py --no-cpu-baseline --no-graph --steps 20 > gpurun_out/abe_$v.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/abe_$v.json')); print('stack=$v eager', round(d['ms_per_step'],3))"
done
