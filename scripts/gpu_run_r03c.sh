#!/bin/bash
# Round-2 pass 29 (1 GPU): 256-token W4A16 tiles on every pair-mode shape again, now that the shared-tile reduction is batched.
O=gpurun_out/r03c; mkdir -p $O
QSPEC_GEMM_NP256=1 timeout 300 python scripts/gpu_ab.py np256_rule > $O/ab.jsonl 2> $O/ab.err
QSPEC_GEMM_NP256=2 timeout 300 python scripts/gpu_ab.py np256_all >> $O/ab.jsonl 2>> $O/ab.err
QSPEC_GEMM_NP256=0 timeout 300 python scripts/gpu_ab.py np128 >> $O/ab.jsonl 2>> $O/ab.err
grep w4a16 $O/ab.jsonl
