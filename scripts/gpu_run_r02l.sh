#!/bin/bash
# Round-2 twelfth pass (1 GPU): grid sweep of the per-rank GEMM shapes of TP = 2/4/8 at the configs[2] token counts.
O=gpurun_out/r02l; mkdir -p $O
timeout 900 python scripts/gpu_gemm_gsweep_tp.py 1,2,4,8 > $O/gsweep_tp.txt 2> $O/gsweep_tp.err; echo "rc=$?" >> $O/gsweep_tp.err
cat $O/gsweep_tp.txt | cut -c1-400
tail -3 $O/gsweep_tp.err
