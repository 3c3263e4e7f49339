#!/bin/bash
# Which resource bounds the up-projection / dgrad-down at C size (16 experts x 8192 rows, d=512, F=2048)?
# streaming vs A-stationary, ring depth, run length, GELU vs ReLU epilogue (tools/gemm_bench.py, isolated kernels).
O=gpurun_out/ares_experiments.txt
: > $O
run() { echo "### $*" >> $O; env "$@" timeout 120 python tools/gemm_bench.py 16 512 2048 8192 >> $O 2>&1; }
run B200MOE_GEMM_ARES=0
run B200MOE_GEMM_ARES=0 GEMM_BENCH_ACT=1
run B200MOE_GEMM_ARES=0 B200MOE_GEMM_PAIR_ALL=1
run B200MOE_GEMM_ARES=0 B200MOE_GEMM_PAIR_ALL=1 GEMM_BENCH_ACT=1
run B200MOE_GEMM_ARES=1
run B200MOE_GEMM_ARES=1 GEMM_BENCH_ACT=1
run B200MOE_GEMM_ARES=1 B200MOE_GEMM_ARES_STAGES=2
run B200MOE_GEMM_ARES=1 B200MOE_GEMM_ARES_STAGES=3
run B200MOE_GEMM_ARES=1 B200MOE_GEMM_ARES_NRUN=2
run B200MOE_GEMM_ARES=1 B200MOE_GEMM_ARES_NRUN=4
echo "### d=256 (8 experts x 16384 rows, F=1024)" >> $O
run2() { echo "### $*" >> $O; env "$@" timeout 120 python tools/gemm_bench.py 8 256 1024 16384 >> $O 2>&1; }
run2 B200MOE_GEMM_ARES=0
run2 B200MOE_GEMM_ARES=0 GEMM_BENCH_ACT=1
run2 B200MOE_GEMM_ARES=1
run2 B200MOE_GEMM_ARES=1 GEMM_BENCH_ACT=1
cat $O
