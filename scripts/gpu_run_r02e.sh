#!/bin/bash
# Round-2 fifth GPU pass: parity suite (seq_attention, rounding-envelope TP checks), mbarrier suspend-hint A/B, bench, 64-seq timeline.
O=gpurun_out/r02e; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -q ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -5 $O/pytest_all.log
L=$PWD/qspec_b200/_C
timeout 300 python scripts/gpu_ab.py default > $O/ab.jsonl 2> $O/ab.err
for ns in 1000 20000 1000000; do
  QSPEC_LIB=$L/libqspec_b200_hint$ns.so timeout 300 python scripts/gpu_ab.py hint$ns >> $O/ab.jsonl 2>> $O/ab.err
done
QSPEC_GEMM_STAGES=8 timeout 300 python scripts/gpu_ab.py stages8 >> $O/ab.jsonl 2>> $O/ab.err
( time timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench.log 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
tail -2 $O/bench.err
timeout 300 python scripts/gpu_timeline.py --seqs 64 > $O/timeline_64.md 2> $O/timeline_64.err
ls -la $O
