#!/bin/bash
# Round-2 final pass b (1 GPU): default bench line with the int8-container GEMM in the per-layer table, ncu launch list of the
# final step and --set full of the draft gate_up GEMM (MODE_A8) and the verify gate_up GEMM (each after the same command exited 0).
O=gpurun_out/r03e; mkdir -p $O
( timeout 900 python bench.py ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
tail -c 400 $O/bench_default.log
CMD="python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline --no-configs1 --no-per-layer --agree-shifts="
$CMD > $O/plain.log 2>&1 &&
ncu --nvtx --nvtx-include "timed/" --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
NCU="ncu --nvtx --nvtx-include timed/ --set full --clock-control none --import-source on"
$NCU -k regex:w4_stream -s 130 -c 1 -o $O/prof_w4a4_c8_m64_gate_up $CMD > $O/ncu_w4a4.log 2>&1
$NCU -k regex:w4_stream -s 386 -c 1 -o $O/prof_w4a16_m256_gate_up $CMD > $O/ncu_w4a16.log 2>&1
ls -la $O
