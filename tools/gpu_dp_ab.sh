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
run default A=1
run gate_nw8 VNB_GATE_ROWS_NW16=0
run gn_nw8 VNB_GN_ROWS_NW32=0
run both_nw8 VNB_GATE_ROWS_NW16=0 VNB_GN_ROWS_NW32=0
run concat_2 VNB_CONCAT_CTAS_PER_SM=2
