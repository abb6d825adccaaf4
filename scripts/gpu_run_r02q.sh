#!/bin/bash
# Round-2 8-GPU pass 2: in-loop kernel timelines (CUPTI through torch.profiler; analysis only) of configs[2] at TP = 8, 4, 2.
O=gpurun_out/r02q; mkdir -p $O
for n in 8 4 2; do
  ( timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29631 scripts/gpu_timeline.py --seqs 64 ) > $O/timeline_64_tp$n.md 2> $O/timeline_64_tp$n.err
  echo "rc=$?" >> $O/timeline_64_tp$n.err
  head -16 $O/timeline_64_tp$n.md | cut -c1-200
done
