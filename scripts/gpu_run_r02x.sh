#!/bin/bash
# Round-2 pass 24 (1 GPU): packed FFMA2 in the had28 mixes -- parity (bit-exact tests), activation-side timings, bench.
O=gpurun_out/r02x; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_full_sizes.py tests/test_gpu_ref_kernels.py tests/test_gpu_model.py -m gpu -q -x ) > $O/pytest.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest.log
tail -4 $O/pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python scripts/gpu_ab_act.py ffma2 > $O/ab_act.jsonl 2> $O/ab_act.err; grep k28 $O/ab_act.jsonl
( timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-per-layer ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.log").read().strip().splitlines()[-1])
print("ms/step", round(d["ms_per_step"],3), "configs1", d["configs1"]["ms_per_step"])
PY
