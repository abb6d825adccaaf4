#!/bin/bash
# Round-2 pass 20 (1 GPU): run-to-run determinism of the 64-sequence step (graph and eager), three embedding shifts.
O=gpurun_out/r02t; mkdir -p $O
timeout 900 python scripts/gpu_determinism.py > $O/determinism.jsonl 2> $O/determinism.err; echo "rc=$?" >> $O/determinism.err
cat $O/determinism.jsonl; tail -3 $O/determinism.err
