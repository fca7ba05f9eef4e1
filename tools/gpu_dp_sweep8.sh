#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
run() {
  local tag=$1; shift
  env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29530 bench.py --gpus $NG --steps 100 --warmup 10 > gpurun_out/dp${NG}_$tag.json 2> gpurun_out/dp${NG}_$tag.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/dp${NG}_$tag.json') if l.startswith('{')][-1])
    print('N=$NG $tag', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s')
except Exception as e:
    print('$tag failed', e); print(open('gpurun_out/dp${NG}_$tag.err').read()[-800:])
PY
}
run b33 VNB_BUCKETS=3,3
run b66 VNB_BUCKETS=6,6
run b66_blocks16 VNB_BUCKETS=6,6 VNB_AR_BLOCKS=16
run b66_blocks32 VNB_BUCKETS=6,6 VNB_AR_BLOCKS=32
run nccl_b66 VNB_ALLREDUCE=nccl VNB_BUCKETS=6,6
run b66_bf16 VNB_BUCKETS=6,6 A=1
