#!/bin/bash
# Round-2 third GPU pass: full parity suite on the CTA-pair kernel, pair A/B microbench, timelines at 4 and 64 sequences.
O=gpurun_out/r02c; mkdir -p $O
( time QSPEC_GEMM_PAIR=1 timeout 900 python -m pytest tests -m gpu -q -x ) > $O/pytest_pair_all.log 2>&1; echo "rc=$?" >> $O/pytest_pair_all.log
tail -5 $O/pytest_pair_all.log
for pair in 0 1; do
  QSPEC_GEMM_PAIR=$pair timeout 600 python scripts/gpu_microbench.py > $O/microbench_pair$pair.json 2> $O/microbench_pair$pair.err
done
for pair in 0 1; do
for spec in "w4a4 64 28672 4096" "w4a16 256 28672 4096" "w4a16 16 28672 4096" "w4a4 4 28672 4096"; do
  echo "==== PAIR=$pair $spec" >> $O/gemm_trace.txt
  QSPEC_GEMM_PAIR=$pair QSPEC_LIB=$PWD/qspec_b200/_C/libqspec_b200_prof.so timeout 120 python scripts/gpu_gemm_trace.py $spec >> $O/gemm_trace.txt 2>&1
done; done
QSPEC_GEMM_PAIR=0 timeout 300 python scripts/gpu_timeline.py --seqs 64 > $O/timeline_64_pair0.md 2> $O/timeline_64_pair0.err
QSPEC_GEMM_PAIR=1 timeout 300 python scripts/gpu_timeline.py --seqs 64 > $O/timeline_64_pair1.md 2> $O/timeline_64_pair1.err
QSPEC_GEMM_PAIR=0 timeout 300 python scripts/gpu_timeline.py --seqs 4 > $O/timeline_4_pair0.md 2> $O/timeline_4_pair0.err
ls -la $O
