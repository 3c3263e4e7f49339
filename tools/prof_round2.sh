#!/bin/bash
# Round-2 ncu captures (run on the GPU box from the repo root; every ncu run follows a plain run of the same command
# that exited 0, B200_PROFILING.md).  Reports land in gpurun_out/; summaries are made on the dev box afterwards.
set -u
O=gpurun_out
NCU="ncu --clock-control none"
# 1. launch list of the default bench command (device time of every launch, cold-cache and serialised: compare shares)
python bench.py --steps 2 --warmup 3 --no-sweep > $O/prof_bench_plain.json 2> $O/prof_bench_plain.err &&
  $NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/r02_launches_bench_A.csv \
    python bench.py --steps 2 --warmup 3 --no-sweep > $O/prof_bench_ncu.log 2>&1
# 2. every kernel of one fwd+bwd step, --set full, at A / 3200 and C / 65533 tokens (second of two steps)
for cfg in "A 8 400" "C 71 923"; do
  set -- $cfg
  python tools/profile_step.py $1 $2 $3 2 > $O/prof_step_$1_plain.log 2>&1 &&
    $NCU --set full -s 18 -c 40 -o $O/r02_ncu_all_$1 -f python tools/profile_step.py $1 $2 $3 2 > $O/prof_step_$1_ncu.log 2>&1
  # the reports are 25-35 MB each and gpurun_out/ travels back only below 64 MiB: summarise on the box, drop the report
  T=$(( $2 * $3 ))
  python tools/ncu_summary.py $O/r02_ncu_all_$1.ncu-rep $O/r02_ncu_all_$1$T.md \
    "ncu --set full, every kernel of one fwd+bwd step, workload $1, $T tokens (tools/profile_step.py $1 $2 $3 2)" > /dev/null
  python tools/ncu_traffic.py $O/r02_ncu_all_$1.ncu-rep $1$T "tools/profile_step.py $1 $2 $3 2" $O/r02_traffic.json
  rm -f $O/r02_ncu_all_$1.ncu-rep
done
# 3. the two epilogue-bound GEMMs with source-level stall samples (C-sized rows: 16 experts x 8192 rows, d=512, F=2048)
python tools/gemm_bench.py 16 512 2048 8192 > $O/prof_gemm_plain.log 2>&1 &&
  $NCU --set full --import-source on -k regex:grouped_gemm -s 3 -c 1 -o $O/r02_ncu_up -f \
    python tools/gemm_bench.py 16 512 2048 8192 > $O/prof_gemm_up.log 2>&1 &&
  $NCU --set full --import-source on -k regex:grouped_gemm -s 27 -c 1 -o $O/r02_ncu_dgrad_down -f \
    python tools/gemm_bench.py 16 512 2048 8192 > $O/prof_gemm_dgd.log 2>&1
python tools/ncu_hot.py $O/r02_ncu_up.ncu-rep grouped_gemm 40 > $O/r02_hot_up.txt 2>&1
python tools/ncu_hot.py $O/r02_ncu_dgrad_down.ncu-rep grouped_gemm 40 > $O/r02_hot_dgrad_down.txt 2>&1
ls -la $O/
