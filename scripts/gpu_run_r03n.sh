#!/bin/bash
# Round-2 last pass (1 GPU): the full GPU suite on the final tree.
O=gpurun_out/r03n; mkdir -p $O
( time timeout 600 python -m pytest tests -m gpu -q -x ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -6 $O/pytest_all.log
