#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=${NG:-2}
run() {
  local tag=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $NG --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/dp_$tag.json 2> gpurun_out/dp_$tag.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/dp_$tag.json') if l.startswith('{')][-1])
    print('$tag', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s')
except Exception as e:
    print('$tag failed', e); print(open('gpurun_out/dp_$tag.err').read()[-800:])
PY
}
run b_6_6 A=1
run b_6_10 VNB_BUCKETS=6,10
run b_10_10 VNB_BUCKETS=10,10
run b_8_9 VNB_BUCKETS=8,9
run b_6_6_b A=1
run b_10_10_b VNB_BUCKETS=10,10
