#!/bin/bash
# Round-2 pass 26 (1 GPU): int8 container of the draft path's activations -- parity tests, model tests, bench A/B.
O=gpurun_out/r02z; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_c8.py tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_tp_onegpu.py -m gpu -q -x ) > $O/pytest.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest.log
tail -30 $O/pytest.log | cut -c1-250
if [ $rc -ne 0 ]; then exit 0; fi
for c in 48 100000; do
  ( QSPEC_C8_MIN_T=$c timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-per-layer --no-configs1 --agree-shifts 19 ) > $O/bench_c8_$c.log 2> $O/bench_c8_$c.err; echo "rc=$?" >> $O/bench_c8_$c.err
  python - <<PY
import json
d=json.loads(open("$O/bench_c8_$c.log").read().strip().splitlines()[-1])
a=d["acceptance"]["agree_by_construction"]
print("QSPEC_C8_MIN_T=$c ms/step", round(d["ms_per_step"],3), "agree", a["draft_acceptance_rate"], round(a["value"],1))
PY
done
