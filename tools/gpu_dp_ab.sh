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
run fuse_148 VNB_WG_FUSE=1
run fuse_100 VNB_WG_FUSE=1 VNB_WG_FUSE_MAXCTA=100
run fuse_64 VNB_WG_FUSE=1 VNB_WG_FUSE_MAXCTA=64
run no_fuse VNB_WG_FUSE=0
for v in 148 100 64; do VNB_WG_FUSE=1 VNB_WG_FUSE_MAXCTA=$v timeout 100 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('single fuse_$v', round(d['ms_per_step'],3))"; done
VNB_WG_FUSE=0 timeout 100 python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('single no_fuse', round(d['ms_per_step'],3))"
