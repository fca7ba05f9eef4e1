#!/bin/bash
# N-GPU scaling of bench.py (run under gpurun --gpus N): N = 1, 2, 4, 8 up to the GPUs visible.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
PREC=${PREC:-bf16x2}
for n in 1 2 4 8; do
  [ $n -gt $NG ] && break
  if [ $n -eq 1 ]; then
    timeout 120 python bench.py --gpus 1 --steps 100 --warmup 10 --no-cpu-baseline --precision $PREC > gpurun_out/scale_${PREC}_$n.json 2> gpurun_out/scale_${PREC}_$n.err
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 100 --warmup 10 --precision $PREC > gpurun_out/scale_${PREC}_$n.json 2> gpurun_out/scale_${PREC}_$n.err
  fi
  echo "N=$n exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/scale_${PREC}_$n.json') if l.startswith('{')][-1])
    print('N=$n', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s', 'e2e', round(d['e2e']['value']))
except Exception as e:
    print('N=$n failed', e); print(open('gpurun_out/scale_${PREC}_$n.err').read()[-1500:])
PY
done
