# usage: bash profiles/final_nN.sh N   -- bench.py under torchrun on N GPUs of one box (the driver's launch line)
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print('N=$N ms', round(d['ms_per_step'],3), 'value', d['value'], 'e2e', round(d['e2e']['ms_per_step'],3), 'ms', d.get('multispeaker',{}).get('ms_per_step'), 'stress', d.get('mas_stress',{}).get('us_per_call'))"
