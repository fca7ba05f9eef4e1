#!/bin/bash
# Round-2 session A: whole GPU parity suite (incl. the new 2-plane mode and the per-mode parity report), bench lines for
# the three operand modes, the reference arm on host CPU and as stock PyTorch on the B200, GEMM microbenchmarks with cuBLAS.
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
run pytest_all 420 python -m pytest tests -m gpu -q -p no:cacheprovider
run bench_bf16 90 python bench.py --steps 100 --warmup 10 --no-cpu-baseline
run bench_bf16x2 90 python bench.py --steps 100 --warmup 10 --precision bf16x2 --no-cpu-baseline
run bench_fp32 90 python bench.py --steps 50 --warmup 10 --precision fp32 --no-cpu-baseline
run gemm_p1 60 python tools/bench_gemm.py r2a_p1
VNB_BENCH_P=2 run gemm_p2 60 python tools/bench_gemm.py r2a_p2
run ref_cuda_fp32 120 python bench.py --impl reference --ref-device cuda --steps 50 --warmup 5
run ref_cuda_autocast 120 python bench.py --impl reference --ref-device cuda --ref-autocast --steps 50 --warmup 5
run ref_cpu 150 python bench.py --impl reference --steps 6 --warmup 2
cat $S
tail -n 30 gpurun_out/pytest_all.log
cat gpurun_out/bench_bf16.log gpurun_out/bench_bf16x2.log gpurun_out/bench_fp32.log gpurun_out/ref_cuda_fp32.log gpurun_out/ref_cuda_autocast.log gpurun_out/ref_cpu.log | cut -c1-400
cat gpurun_out/gemm_p1.log gpurun_out/gemm_p2.log
