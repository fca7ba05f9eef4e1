#!/bin/bash
# Round-2 final session: full GPU suite, smoke, the bench lines of every mode and workload, the reference arm,
# the ncu launch list of the headline command and one --set full capture of the dominant GEMM.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S=gpurun_out/shot_status.txt
: > $S
DEADLINE=$(( SECONDS + ${SHOT_SECONDS:-1500} ))
run() {
    local name=$1 t=$2; shift 2
    local left=$(( DEADLINE - SECONDS ))
    if [ $left -lt 15 ]; then echo "$name skipped (deadline)" >> $S; return; fi
    [ $t -gt $left ] && t=$left
    local t0=$SECONDS
    timeout "$t" "$@" > gpurun_out/$name.log 2> gpurun_out/$name.err
    echo "$name exit $? $(( SECONDS - t0 ))s" >> $S
}
run pytest_all 400 python -m pytest tests -m gpu -x -q -p no:cacheprovider
run smoke 100 python __graft_entry__.py smoke
run bench_default 300 python bench.py
run bench_bf16 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --precision bf16
run bench_fp32 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --precision fp32
run bench_reference 300 python bench.py --impl reference --steps 3 --warmup 1
run wl_yaml 120 python bench.py --workload vcc2020_yaml --steps 30 --warmup 5 --no-cpu-baseline
run wl_aishell3 100 python bench.py --workload aishell3 --steps 50 --warmup 10 --no-cpu-baseline
run wl_cdvae 100 python bench.py --workload cdvae --steps 50 --warmup 10 --no-cpu-baseline
run wl_decode 100 python bench.py --workload decode --steps 50 --warmup 10 --no-cpu-baseline
run gemm_p2 100 env VNB_BENCH_P=2 python tools/bench_gemm.py r2final
run kernel_timeline 100 python tools/diag_kernel_timeline.py bf16x2
if grep -q "bench_default exit 0" $S; then
run ncu_launches 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
fi
if grep -q "gemm_p2 exit 0" $S; then
run ncu_full 200 env VNB_BENCH_P=2 ncu --set full --clock-control none --import-source on --kernel-id ::regex:conv_fwd_tc:3 -c 1 -o gpurun_out/conv_fwd_dominant_r2 -f python tools/bench_gemm.py ncu 0
fi
cat $S
tail -n 5 gpurun_out/pytest_all.log
tail -n 2 gpurun_out/smoke.log
for f in bench_default bench_bf16 bench_fp32 bench_reference wl_yaml wl_aishell3 wl_cdvae wl_decode; do python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/$f.log') if l.startswith('{')][-1])
    print('$f', d.get('ms_per_step'), d.get('value'), d.get('unit'), 'e2e', (d.get('e2e') or {}).get('value'), 'launches', d.get('gpu_launches'),
          'roof', (d.get('roofline') or {}).get('frac'), 'cpu', (d.get('cpu_baseline') or {}).get('value'))
except Exception as e:
    print('$f failed', e); print(open('gpurun_out/$f.err').read()[-600:])
PY
done
