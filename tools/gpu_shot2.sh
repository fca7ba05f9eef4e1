#!/bin/bash
# Bounded GPU session: full parity suite on the current defaults, bench lines (bf16, fp32-faithful, reference arm),
# microbenchmarks, ncu launch list of the bench command and one --set full capture of each second-generation kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S=gpurun_out/shot_status.txt
: > $S
DEADLINE=$(( SECONDS + ${SHOT_SECONDS:-400} ))
run() {
    local name=$1 t=$2; shift 2
    local left=$(( DEADLINE - SECONDS ))
    if [ $left -lt 15 ]; then echo "$name skipped (deadline)" >> $S; return; fi
    [ $t -gt $left ] && t=$left
    local t0=$SECONDS
    timeout "$t" "$@" > gpurun_out/$name.log 2> gpurun_out/$name.err
    echo "$name exit $? $(( SECONDS - t0 ))s" >> $S
}
run pytest_all 200 python -m pytest tests -m gpu -q -p no:cacheprovider
run bench_bf16 100 python bench.py --steps 200 --warmup 10
run bench_default 100 python bench.py
run bench_fp32 100 python bench.py --steps 100 --warmup 10 --precision fp32 --no-cpu-baseline
run kernels 80 python tools/bench_kernels.py --mem-only --tag final
[ -n "$SKIP_NCU" ] || run ncu_launches 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
[ -n "$SKIP_NCU" ] || run ncu_wn 100 ncu --set full --clock-control none --import-source on --kernel-id ::regex:wn_:3 -o gpurun_out/wn_v2 -f python tools/bench_kernels.py --mem-only --tag ncu
run bench_ref 100 python bench.py --impl reference --steps 6 --warmup 2
run smoke 100 python __graft_entry__.py smoke
cat $S
tail -n 4 gpurun_out/pytest_all.log
cat gpurun_out/bench_bf16.log gpurun_out/bench_default.log gpurun_out/smoke.log
grep wn_ gpurun_out/kernels.log
