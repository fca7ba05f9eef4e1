#!/bin/bash
# One bounded GPU session: parity tests of the new paths with the second-generation streaming kernels switched on,
# the bench line with both kernel generations, the memory-bound microbenchmarks, the conversion throughput, then
# the rest of the GPU suite.  Every step has its own timeout inside one overall deadline and writes under
# gpurun_out/; a step that fails does not stop the ones after it.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S=gpurun_out/shot_status.txt
: > $S
DEADLINE=$(( SECONDS + ${SHOT_SECONDS:-540} ))
run() {  # name, timeout, command...
    local name=$1 t=$2; shift 2
    local left=$(( DEADLINE - SECONDS ))
    if [ $left -lt 20 ]; then echo "$name skipped (deadline)" >> $S; return; fi
    [ $t -gt $left ] && t=$left
    local t0=$SECONDS
    timeout "$t" "$@" > gpurun_out/$name.log 2> gpurun_out/$name.err
    echo "$name exit $? $(( SECONDS - t0 ))s" >> $S
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader >> $S 2>&1
export VNB_EW_V2=1 VNB_FAST_GATE=1
run pytest_new 240 python -m pytest tests/test_gpu_convert.py tests/test_gpu_ew_v2.py -q -p no:cacheprovider
run bench_v2 100 python bench.py --steps 20 --warmup 5
export VNB_EW_V2=0 VNB_FAST_GATE=0
run bench_v1 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline
run kernels_v1 80 python tools/bench_kernels.py --mem-only --tag v1
export VNB_EW_V2=1 VNB_FAST_GATE=1
run kernels_v2 80 python tools/bench_kernels.py --mem-only --tag v2
run convert 120 python tools/bench_convert.py --out gpurun_out/convert.json
run pytest_old_v2 400 python -m pytest tests -m gpu -q -p no:cacheprovider --ignore=tests/test_gpu_convert.py --ignore=tests/test_gpu_ew_v2.py
cat $S
tail -4 gpurun_out/pytest_new.log gpurun_out/pytest_old_v2.log
cat gpurun_out/bench_v2.log gpurun_out/bench_v1.log
