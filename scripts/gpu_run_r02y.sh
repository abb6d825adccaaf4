#!/bin/bash
# Round-2 pass 25 (1 GPU): W4A8 (int8 activations as TMA delivers them) next to W4A4 on the current kernels, M in {4, 16, 64}.
O=gpurun_out/r02y; mkdir -p $O
timeout 600 python scripts/gpu_microbench.py --w4a8 > $O/microbench_w4a8.jsonl 2> $O/microbench_w4a8.err; echo "rc=$?" >> $O/microbench_w4a8.err
python - <<PY
import json
for l in open("$O/microbench_w4a8.jsonl"):
    r=json.loads(l); print(r["kernel"], r["shape"], r["M"], r["us_avg"])
PY
