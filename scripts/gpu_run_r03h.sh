#!/bin/bash
# Round-2 pass 34 (1 GPU): FMA form of the had108 / had40 mixes -- Qwen parity tests (bit-exact kernels, one-layer decode vs oracle), step time.
O=gpurun_out/r03h; mkdir -p $O
( timeout 600 python -m pytest tests/test_gpu_qwen.py -m gpu -q -x ) > $O/pytest_qwen.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest_qwen.log
tail -4 $O/pytest_qwen.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 600 python scripts/gpu_qwen_step.py > $O/qwen_step.jsonl 2> $O/qwen_step.err; echo "rc=$?" >> $O/qwen_step.err
cat $O/qwen_step.jsonl; tail -2 $O/qwen_step.err
