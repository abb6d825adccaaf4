#!/bin/bash
# Round-2 pass 15 (1 GPU): allocation-free arg max alone in a fresh process, then the whole GPU suite (it runs last), then bench.
O=gpurun_out/r02o; mkdir -p $O
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x ) > $O/pytest_no_alloc_alone.log 2>&1; echo "rc=$?" >> $O/pytest_no_alloc_alone.log
tail -4 $O/pytest_no_alloc_alone.log
( time timeout 1200 python -m pytest tests -m gpu -q -x ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -6 $O/pytest_all.log
( timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
tail -c 1500 $O/bench_default.log
