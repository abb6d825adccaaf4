#!/bin/bash
# Round-2 8-GPU pass 3: A/B of the two-step fused all-reduce on configs[2] at TP = 8 and 4 (QSPEC_TP_SCATTER_MIN: 100000 = one-step
# everywhere, 1000 = two-step on the 256-token verify exchanges of TP = 8 only, 384 = default).
O=gpurun_out/r02s; mkdir -p $O
run() {  # name, nproc, scatter_min
  ( QSPEC_TP_SCATTER_MIN=$3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29651 bench.py --gpus $2 --steps 20 --warmup 3 --no-per-layer --agree-shifts 19 ) > $O/bench_$1.log 2> $O/bench_$1.err
  echo "rc=$?" >> $O/bench_$1.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$1.log").read().strip().splitlines()[-1])
    print("$1: ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1))
except Exception as e:
    print("$1: no line", e)
PY
}
run n8_onestep 8 100000
run n8_verify 8 1000
run n8_default 8 384
run n4_onestep 4 100000
run n4_default 4 384
run n8_all 8 1
