#!/bin/bash
# Round-2 fourth GPU pass: parity suite; same-box A/B of the round-1 GEMM, the current one (pairs off / on / from M = 128), and the watchdog build.
O=gpurun_out/r02d; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -q ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -5 $O/pytest_all.log
L=$PWD/qspec_b200/_C
QSPEC_LIB=$PWD/build/libqspec_b200_r01gemm.so timeout 300 python scripts/gpu_ab.py r01gemm > $O/ab.jsonl 2> $O/ab.err
QSPEC_GEMM_PAIR=0 timeout 300 python scripts/gpu_ab.py cur_pair0 >> $O/ab.jsonl 2>> $O/ab.err
QSPEC_GEMM_PAIR=0 QSPEC_LIB=$L/libqspec_b200_wd.so timeout 300 python scripts/gpu_ab.py wd_pair0 >> $O/ab.jsonl 2>> $O/ab.err
QSPEC_GEMM_PAIR=1 timeout 300 python scripts/gpu_ab.py cur_pair1 >> $O/ab.jsonl 2>> $O/ab.err
QSPEC_LIB=$PWD/build/libqspec_b200_r01gemm.so timeout 300 python scripts/gpu_ab.py r01gemm_again >> $O/ab.jsonl 2>> $O/ab.err
ls -la $O
