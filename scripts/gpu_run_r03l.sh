#!/bin/bash
# Round-2 pass 38 (1 GPU): ncu --set full of the had108 rotation kernel in the Qwen variant's step (2 layers, 4 sequences), after the
# same command exited 0 without ncu.
O=gpurun_out/r03l; mkdir -p $O
CMD="python scripts/gpu_qwen_step.py --layers 2 --seqs 4 --steps 2 --warmup 1"
$CMD > $O/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:hadamard_k108 -s 2 -c 1 -o $O/prof_k108_m4 $CMD > $O/ncu_k108.log 2>&1
tail -3 $O/ncu_k108.log; ls -la $O
