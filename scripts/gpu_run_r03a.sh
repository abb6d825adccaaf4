#!/bin/bash
# Round-2 pass 27 (2 GPUs): int8 container under tensor parallelism -- emulation test, the 2-GPU protocol test, configs[2] at TP = 2 A/B.
O=gpurun_out/r03a; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_c8.py tests/test_gpu_tp.py -m gpu -q -x ) > $O/pytest.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest.log
tail -6 $O/pytest.log | cut -c1-250
if [ $rc -ne 0 ]; then exit 0; fi
for c in 48 100000; do
  ( QSPEC_C8_MIN_T=$c timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29661 bench.py --gpus 2 --steps 20 --warmup 3 --no-per-layer --agree-shifts 19 ) > $O/bench_n2_c8_$c.log 2> $O/bench_n2_c8_$c.err; echo "rc=$?" >> $O/bench_n2_c8_$c.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_n2_c8_$c.log").read().strip().splitlines()[-1])
    a=d["acceptance"]["agree_by_construction"]
    print("N=2 QSPEC_C8_MIN_T=$c ms/step", round(d["ms_per_step"],3), "agree", a["draft_acceptance_rate"], round(a["value"],1))
except Exception as e:
    print("no line", e)
PY
done
