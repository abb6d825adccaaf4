#!/bin/bash
# Round-2 second GPU pass: drop-in test, CTA-pair (cta_group::2) GEMM parity + A/B timing, bench line, role timelines.
O=gpurun_out/r02b; mkdir -p $O
( QSPEC_GEMM_PAIR=0 timeout 600 python -m pytest tests/test_gpu_dropin_reference_model.py -m gpu -q ) > $O/pytest_dropin.log 2>&1; echo "rc=$?" >> $O/pytest_dropin.log
tail -3 $O/pytest_dropin.log
( time QSPEC_GEMM_PAIR=1 timeout 400 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_w4a8.py tests/test_gpu_full_sizes.py -m gpu -q -x ) > $O/pytest_pair_gemm.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest_pair_gemm.log
tail -8 $O/pytest_pair_gemm.log
if [ $rc -eq 0 ]; then
  ( time QSPEC_GEMM_PAIR=1 timeout 900 python -m pytest tests -m gpu -q ) > $O/pytest_pair_all.log 2>&1; echo "rc=$?" >> $O/pytest_pair_all.log
  tail -5 $O/pytest_pair_all.log
  for pair in 0 1; do
    QSPEC_GEMM_PAIR=$pair timeout 600 python scripts/gpu_microbench.py --quick > $O/microbench_pair$pair.json 2> $O/microbench_pair$pair.err
  done
  ( time QSPEC_GEMM_PAIR=1 timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench_pair1.log 2> $O/bench_pair1.err; echo "bench rc=$?" >> $O/bench_pair1.err
  tail -3 $O/bench_pair1.err
fi
( time QSPEC_GEMM_PAIR=0 timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench_pair0.log 2> $O/bench_pair0.err; echo "bench rc=$?" >> $O/bench_pair0.err
tail -3 $O/bench_pair0.err
for pair in 0 1; do
for spec in "w4a4 64 28672 4096" "w4a16 256 28672 4096" "w4a16 16 28672 4096" "w4a16 16 4096 4096"; do
  echo "==== PAIR=$pair $spec" >> $O/gemm_trace.txt
  QSPEC_GEMM_PAIR=$pair QSPEC_LIB=$PWD/qspec_b200/_C/libqspec_b200_prof.so timeout 120 python scripts/gpu_gemm_trace.py $spec >> $O/gemm_trace.txt 2>&1
done; done
ls -la $O
