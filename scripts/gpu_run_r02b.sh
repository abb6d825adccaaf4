#!/bin/bash
# Round-2 second GPU pass: drop-in test, bench line, GEMM role timelines at M = 64 / 256.
O=gpurun_out/r02b; mkdir -p $O
( time timeout 600 python -m pytest tests/test_gpu_dropin_reference_model.py -m gpu -q ) > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
( time timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench.log 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
tail -c 400 $O/bench.log; tail -3 $O/bench.err
for spec in "w4a4 64 28672 4096" "w4a4 64 4096 14336" "w4a16 256 28672 4096" "w4a16 256 4096 4096" "w4a16 16 28672 4096"; do
  echo "==== $spec" >> $O/gemm_trace.txt
  QSPEC_LIB=$PWD/qspec_b200/_C/libqspec_b200_prof.so timeout 120 python scripts/gpu_gemm_trace.py $spec >> $O/gemm_trace.txt 2>&1
done
ls -la $O
