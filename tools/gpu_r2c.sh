#!/bin/bash
# Round-2 session C: workload tests + bench lines of every named workload.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S=gpurun_out/shot_status.txt
: > $S
DEADLINE=$(( SECONDS + ${SHOT_SECONDS:-900} ))
run() {
    local name=$1 t=$2; shift 2
    local left=$(( DEADLINE - SECONDS ))
    if [ $left -lt 15 ]; then echo "$name skipped (deadline)" >> $S; return; fi
    [ $t -gt $left ] && t=$left
    local t0=$SECONDS
    timeout "$t" "$@" > gpurun_out/$name.log 2> gpurun_out/$name.err
    echo "$name exit $? $(( SECONDS - t0 ))s" >> $S
}
run pytest_r2 300 python -m pytest tests/test_gpu_round2.py tests/test_gpu_parity_modes.py -q -p no:cacheprovider
run wl_vcc2020 100 python bench.py --steps 100 --warmup 10 --no-cpu-baseline
run wl_yaml 150 python bench.py --workload vcc2020_yaml --steps 20 --warmup 5 --no-cpu-baseline
run wl_yaml_bf16 150 python bench.py --workload vcc2020_yaml --steps 20 --warmup 5 --no-cpu-baseline --precision bf16
run wl_aishell3 100 python bench.py --workload aishell3 --steps 50 --warmup 10 --no-cpu-baseline
run wl_cdvae 100 python bench.py --workload cdvae --steps 50 --warmup 10 --no-cpu-baseline
run wl_decode 100 python bench.py --workload decode --steps 50 --warmup 10 --no-cpu-baseline
run ref_decode_cuda 100 python bench.py --impl reference --workload decode --ref-device cuda --steps 20 --warmup 3
run ref_yaml_cuda 100 python bench.py --impl reference --workload vcc2020_yaml --ref-device cuda --steps 10 --warmup 3
run ref_aishell_cuda 100 python bench.py --impl reference --workload aishell3 --ref-device cuda --steps 20 --warmup 3
cat $S
tail -n 15 gpurun_out/pytest_r2.log
for f in wl_vcc2020 wl_yaml wl_yaml_bf16 wl_aishell3 wl_cdvae wl_decode ref_decode_cuda ref_yaml_cuda ref_aishell_cuda; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/$f.log').read().strip().splitlines()[-1])
    r=d.get('roofline',{})
    print('$f', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s e2e',round(d['e2e']['value']),'launches/step',d['gpu_launches']//max(1,d['steps']),'graph',d['config'].get('cuda_graph'),'step_frac',round(d['config'].get('step_roofline_frac_tensor',0),3),'kernel frac',round(r.get('frac',0),3))
except Exception as e:
    print('$f failed', e); print(open('gpurun_out/$f.err').read()[-1200:])
PY
done
