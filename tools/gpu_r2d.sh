#!/bin/bash
# Round-2 session D: full GPU suite on the batched conditioning path, bench, ncu launch list of the headline mode.
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
run pytest_all 400 python -m pytest tests -m gpu -q -p no:cacheprovider
run bench_bf16x2 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline
run bench_bf16 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --precision bf16
run wl_cdvae 100 python bench.py --workload cdvae --steps 50 --warmup 10 --no-cpu-baseline
run smoke 100 python __graft_entry__.py smoke
if grep -q "bench_bf16x2 exit 0" $S; then
run ncu_launches 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bf16x2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
fi
cat $S
tail -n 12 gpurun_out/pytest_all.log
for f in bench_bf16x2 bench_bf16 wl_cdvae; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/$f.log').read().strip().splitlines()[-1])
    r=d.get('roofline',{})
    print('$f', round(d['ms_per_step'],3),'ms', round(d['value']),'f/s e2e',round(d['e2e']['value']),'launches/step',d['gpu_launches']//max(1,d['steps']),'graph',d['config'].get('cuda_graph'),'step_frac',round(d['config'].get('step_roofline_frac_tensor',0),3),'kernel frac',round(r.get('frac',0),3))
except Exception as e:
    print('$f failed', e); print(open('gpurun_out/$f.err').read()[-1200:])
PY
done
tail -3 gpurun_out/smoke.log
python tools/agg_launches.py gpurun_out/launches_bf16x2.csv 2>/dev/null | head -60
