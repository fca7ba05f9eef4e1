#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=${NG:-2}
run() {
  local tag=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $NG --steps 100 --warmup 10 > gpurun_out/dp_$tag.json 2> gpurun_out/dp_$tag.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/dp_$tag.json') if l.startswith('{')][-1])
    print('$tag', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s')
except Exception as e:
    print('$tag failed', e); print(open('gpurun_out/dp_$tag.err').read()[-800:])
PY
}
run default A=1
run no_overlap VNB_NO_OVERLAP=1
run buckets_1_1 VNB_BUCKETS=1,1
run buckets_2_1 VNB_BUCKETS=2,1
run buckets_6_6 VNB_BUCKETS=6,6
run blocks_32 VNB_AR_BLOCKS=32
run blocks_64 VNB_AR_BLOCKS=64
run blocks_8 VNB_AR_BLOCKS=8
run nccl VNB_ALLREDUCE=nccl
run nccl_no_overlap VNB_ALLREDUCE=nccl VNB_NO_OVERLAP=1
