#!/bin/bash
# Round-2 pass 35 (1 GPU): in-loop timeline of the Qwen2.5-14B variant at 4 sequences (CUPTI through torch.profiler; analysis only).
O=gpurun_out/r03i; mkdir -p $O
timeout 300 python scripts/gpu_timeline.py --model qwen14b --seqs 4 > $O/timeline_qwen_4.md 2> $O/timeline_qwen_4.err; echo "rc=$?" >> $O/timeline_qwen_4.err
head -24 $O/timeline_qwen_4.md | cut -c1-200; grep -n "draft pass" -A14 $O/timeline_qwen_4.md | cut -c1-130
