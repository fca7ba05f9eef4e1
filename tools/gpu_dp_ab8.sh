#!/bin/bash
# 8-GPU A/B of the all-reduce slicing (run under gpurun --gpus 8)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
run() {
  local tag=$1; shift
  env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29530 bench.py --gpus $NG --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/dp${NG}_$tag.json 2> gpurun_out/dp${NG}_$tag.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/dp${NG}_$tag.json') if l.startswith('{')][-1])
    print('N=$NG $tag', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s')
except Exception as e:
    print('$tag failed', e); print(open('gpurun_out/dp${NG}_$tag.err').read()[-800:])
PY
}
timeout 100 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', round(d['ms_per_step'],3))"
run b_6_6 A=1
run b_10_10 VNB_BUCKETS=10,10
run b_6_6_b A=1
run b_10_10_b VNB_BUCKETS=10,10
run b_6_6_late_cond VNB_EARLY_COND=0
