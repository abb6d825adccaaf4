#!/bin/bash
# Round-2 pass 33 (1 GPU): supplementary line -- speculative step of the Qwen2.5-14B variant at 64 and 4 sequences.
O=gpurun_out/r03g; mkdir -p $O
timeout 600 python scripts/gpu_qwen_step.py > $O/qwen_step.jsonl 2> $O/qwen_step.err; echo "rc=$?" >> $O/qwen_step.err
cat $O/qwen_step.jsonl; tail -3 $O/qwen_step.err
