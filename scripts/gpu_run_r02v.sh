#!/bin/bash
# Round-2 pass 22 (1 GPU): engine stream ordering fix -- determinism probe again, the model tests, bench.
O=gpurun_out/r02v; mkdir -p $O
timeout 900 python scripts/gpu_determinism.py > $O/determinism.jsonl 2> $O/determinism.err; echo "rc=$?" >> $O/determinism.err
cat $O/determinism.jsonl; tail -2 $O/determinism.err
( timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_tp_onegpu.py -m gpu -q -x ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
( timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.log").read().strip().splitlines()[-1])
print("ms/step", round(d["ms_per_step"],3), "configs1", d["configs1"]["ms_per_step"], d["acceptance"]["agree_by_construction"]["tried"])
PY
