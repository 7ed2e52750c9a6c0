# same-box A/B of two builds of the library: profiles/_ab (the committed HEAD, built by hand) against glowtts_b200/lib
for v in head work head work; do
  if [ $v = head ]; then export GTK_LIB=$PWD/profiles/_ab/libglowtts_b200.so; else unset GTK_LIB; fi
  timeout 300 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/abl_$v.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/abl_$v.json')); print('$v graph', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])"
done
