timeout 600 python -m pytest tests/test_gpu_stack_r2.py tests/test_gpu_parity_r2.py tests/test_gpu_prosody_r2.py -x -q -m gpu 2>&1 | tail -5
for v in 0 1 0 1; do
  GTK_WB_POST=$v timeout 300 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/abp_$v.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/abp_$v.json')); print('post=$v graph', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])"
done
