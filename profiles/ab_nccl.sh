# N=2 sweep of NCCL knobs for the overlapped gradient all-reduce (bench.py under torchrun, 20 steps each)
mkdir -p gpurun_out
run() {
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline --steps 20 > gpurun_out/abn.json 2> gpurun_out/abn.err
  python -c "
import json; d=json.load(open('gpurun_out/abn.json')); print('$*: ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3))" || tail -3 gpurun_out/abn.err
}
run X=0
run NCCL_MAX_CTAS=16
run NCCL_NTHREADS=256
run NCCL_PROTO=Simple
run NCCL_PROTO=LL128
run NCCL_ALGO=Tree
run X=0
