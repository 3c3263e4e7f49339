#!/bin/bash
O=gpurun_out/gemm_experiments2.txt
: > $O
run() { echo "### $*" >> $O; env "$@" timeout 120 python tools/gemm_bench.py 16 512 2048 8192 >> $O 2>&1; }
run X=0
run B200MOE_GEMM_PAIR_ALL=1
run B200MOE_GEMM_PAIR_ALL=1 GEMM_BENCH_ACT=1
run2() { echo "### d=256: $*" >> $O; env "$@" timeout 120 python tools/gemm_bench.py 8 256 1024 16384 >> $O 2>&1; }
run2 X=0
run2 B200MOE_GEMM_PAIR_ALL=1
run3() { echo "### d=1024: $*" >> $O; env "$@" timeout 120 python tools/gemm_bench.py 16 1024 4096 2048 >> $O 2>&1; }
run3 X=0
cat $O
