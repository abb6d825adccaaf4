#!/bin/bash
# Round-2 pass 16 (1 GPU): decode-sized W4A4 after splitting the activation-widening paths: A/B table, GEMM parity, bench.
O=gpurun_out/r02p; mkdir -p $O
timeout 300 python scripts/gpu_ab.py split > $O/ab.jsonl 2> $O/ab.err
cat $O/ab.jsonl
( timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_full_sizes.py tests/test_gpu_w4a8.py -m gpu -q -x ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
( timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.log").read().strip().splitlines()[-1])
print("ms/step", round(d["ms_per_step"],3), "configs1", d["configs1"]["ms_per_step"])
for w in ("workload","configs1"):
    print({k:(v["w4a4"]["us"], v["w4a16"]["us"]) for k,v in d["per_layer"][w].items()})
PY
