#!/bin/bash
# Round-2 session B: third-generation conv GEMM (CTA pairs, multi-plane stages) - parity, then timing.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S=gpurun_out/shot_status.txt
: > $S
DEADLINE=$(( SECONDS + ${SHOT_SECONDS:-600} ))
run() {
    local name=$1 t=$2; shift 2
    local left=$(( DEADLINE - SECONDS ))
    if [ $left -lt 15 ]; then echo "$name skipped (deadline)" >> $S; return; fi
    [ $t -gt $left ] && t=$left
    local t0=$SECONDS
    timeout "$t" "$@" > gpurun_out/$name.log 2> gpurun_out/$name.err
    echo "$name exit $? $(( SECONDS - t0 ))s" >> $S
}
export VNB_GEMM_V3=2
VNB_TC3_PAIR=0 run tc3_single 150 python -m pytest tests/test_gpu_tc.py tests/test_gpu_round2.py -q -p no:cacheprovider -k "conv_fwd or data_gradient or groupnorm or workload" -x
VNB_TC3_PAIR=1 run tc3_pair 150 python -m pytest tests/test_gpu_tc.py tests/test_gpu_round2.py -q -p no:cacheprovider -k "conv_fwd or data_gradient or groupnorm or workload" -x
unset VNB_GEMM_V3
run pytest_all 300 python -m pytest tests -m gpu -q -p no:cacheprovider
run bench_bf16 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline
run bench_bf16x2 90 python bench.py --steps 100 --warmup 10 --precision bf16x2 --no-cpu-baseline
run bench_fp32 90 python bench.py --steps 50 --warmup 10 --precision fp32 --no-cpu-baseline
VNB_GEMM_V3=0 run bench_bf16_v2 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline
cat $S
tail -n 12 gpurun_out/tc3_single.log; tail -n 12 gpurun_out/tc3_pair.log; tail -n 12 gpurun_out/pytest_all.log
for f in bench_bf16 bench_bf16x2 bench_fp32 bench_bf16_v2; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/$f.log').read().strip().splitlines()[-1])
    print('$f', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s e2e',round(d['e2e']['value']),'launches/step',d['gpu_launches']//d['steps'],'roofline',round(d['roofline']['frac'],3),round(d['roofline']['kernel_ms']*1e3,2),'us')
except Exception as e:
    print('$f failed', e); print(open('gpurun_out/$f.err').read()[-1500:])
PY
done
