#!/bin/bash
# Round-2 pass 23 (1 GPU): full GPU suite, configs[4] sweep from HEAD, default bench (+ reference arm), ncu launch list and
# --set full captures of the dominant kernels of the 64-sequence step (each after the same command exited 0 without ncu).
O=gpurun_out/r02w; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -6 $O/pytest_all.log
timeout 900 python scripts/gpu_microbench.py > $O/microbench.jsonl 2> $O/microbench.err; echo "rc=$?" >> $O/microbench.err
( timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
( timeout 600 python bench.py --impl reference --steps 2 --warmup 1 ) > $O/bench_reference.log 2> $O/bench_reference.err; echo "rc=$?" >> $O/bench_reference.err
tail -c 600 $O/bench_reference.log
CMD="python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline --no-configs1 --no-per-layer --agree-shifts="
$CMD > $O/plain.log 2>&1 &&
ncu --nvtx --nvtx-include "timed/" --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
NCU="ncu --nvtx --nvtx-include timed/ --set full --clock-control none --import-source on"
$NCU -k regex:w4_stream -s 386 -c 1 -o $O/prof_w4a16_m256_gate_up $CMD > $O/ncu_w4a16.log 2>&1
$NCU -k regex:w4_stream -s 2 -c 1 -o $O/prof_w4a4_m64_gate_up $CMD > $O/ncu_w4a4.log 2>&1
$NCU -k regex:hadamard_k28 -s 1 -c 1 -o $O/prof_k28_m64 $CMD > $O/ncu_k28_m64.log 2>&1
$NCU -k regex:hadamard_k28 -s 97 -c 1 -o $O/prof_k28_m256 $CMD > $O/ncu_k28_m256.log 2>&1
$NCU -k regex:seq_attention -s 97 -c 1 -o $O/prof_seq_attention_m256 $CMD > $O/ncu_seqattn.log 2>&1
$NCU -k regex:norm_kernel -s 3 -c 1 -o $O/prof_norm_quant_m64 $CMD > $O/ncu_norm.log 2>&1
ls -la $O
