#!/bin/bash
# Round-2 first GPU pass: parity tests, bench line, UMMA issue-rate probe, GEMM role timelines, configs[4] sweep.
O=gpurun_out/r02a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
( time timeout 900 python -m pytest tests -m gpu -q ) > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
( time timeout 600 python bench.py --steps 20 --warmup 3 ) > $O/bench.log 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
tail -c 600 $O/bench.log; tail -3 $O/bench.err
timeout 120 ./build/umma_probe > $O/umma_probe.txt 2>&1; echo "probe rc=$?" >> $O/umma_probe.txt
for spec in "w4a4 64 28672 4096" "w4a4 64 4096 14336" "w4a16 256 28672 4096" "w4a16 256 4096 4096" "w4a16 16 28672 4096"; do
  echo "==== $spec" >> $O/gemm_trace.txt
  QSPEC_LIB=$PWD/qspec_b200/_C/libqspec_b200_prof.so timeout 120 python scripts/gpu_gemm_trace.py $spec >> $O/gemm_trace.txt 2>&1
done
timeout 600 python scripts/gpu_microbench.py > $O/microbench.json 2> $O/microbench.err; echo "microbench rc=$?" >> $O/microbench.err
ls -la $O
