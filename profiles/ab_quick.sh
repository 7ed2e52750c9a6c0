timeout 900 python -m pytest tests -x -q -m gpu -k "wgrad or prosody or parity or stack or ddp or overlap" 2>&1 | tail -3
for v in 1 2; do
  timeout 300 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/abq_$v.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/abq_$v.json')); print('run $v graph', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])"
done
