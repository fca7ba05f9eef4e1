#!/bin/bash
# 8-GPU check of the high-priority communication stream: full scaling run with the defaults, then N=8 at normal priority.
cd "$(dirname "$0")/.."
bash tools/gpu_scale.sh
NG=$(nvidia-smi -L | wc -l)
VNB_COMM_PRIO=0 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NG --steps 100 --warmup 10 > gpurun_out/scale_prio0_$NG.json 2> gpurun_out/scale_prio0_$NG.err
python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/scale_prio0_$NG.json') if l.startswith('{')][-1])
    print('N=$NG normal priority', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s')
except Exception as e:
    print('failed', e); print(open('gpurun_out/scale_prio0_$NG.err').read()[-800:])
PY
