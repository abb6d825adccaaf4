#!/bin/bash
# Round-2 pass 18 (2 GPUs): two-step fused all-reduce -- one-GPU emulation tests, the 2-GPU protocol test in both forms,
# configs[2] at TP = 2 with the two-step form forced on the verify pass (A/B against the default).
O=gpurun_out/r02r; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_tp_onegpu.py tests/test_gpu_tp.py tests/test_gpu_kernels.py -m gpu -q -x -k "tp" ) > $O/pytest_tp.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest_tp.log
tail -25 $O/pytest_tp.log | cut -c1-300
if [ $rc -ne 0 ]; then exit 0; fi
for sm in 384 200; do
  ( QSPEC_TP_SCATTER_MIN=$sm timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 2 --steps 20 --warmup 3 --no-per-layer --agree-shifts 19 ) > $O/bench_n2_sm$sm.log 2> $O/bench_n2_sm$sm.err; echo "rc=$?" >> $O/bench_n2_sm$sm.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_n2_sm$sm.log").read().strip().splitlines()[-1])
    print("scatter_min $sm: ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1))
except Exception as e:
    print("scatter_min $sm: no line", e)
PY
done
