#!/bin/bash
# Round-2 pass 21 (1 GPU): bisect the run-to-run difference found by scripts/gpu_determinism.py.
O=gpurun_out/r02u; mkdir -p $O
timeout 900 python scripts/gpu_determinism_bisect.py > $O/bisect.jsonl 2> $O/bisect.err; echo "rc=$?" >> $O/bisect.err
cat $O/bisect.jsonl; tail -5 $O/bisect.err
