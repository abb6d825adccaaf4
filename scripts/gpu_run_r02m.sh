#!/bin/bash
# Round-2 pass 13 (1 GPU): batched shared-tile reduction + peer-memory arg max (one-GPU emulation): parity, sweep, bench.
O=gpurun_out/r02m; mkdir -p $O
( timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_full_sizes.py tests/test_gpu_tp_onegpu.py tests/test_gpu_w4a8.py -m gpu -q -x ) > $O/pytest.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest.log
tail -5 $O/pytest.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 900 python scripts/gpu_gemm_gsweep_tp.py 1,2,4,8 > $O/gsweep_tp.txt 2> $O/gsweep_tp.err; echo "rc=$?" >> $O/gsweep_tp.err
cut -c1-330 $O/gsweep_tp.txt
( timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
QSPEC_SPLIT_FUSED=0 timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs1 --no-per-layer > $O/bench_nosplit.log 2> $O/bench_nosplit.err
python - <<PY
import json
for n in ("default", "nosplit"):
    try:
        d=json.loads(open("$O/bench_%s.log" % n).read().strip().splitlines()[-1])
        print(n, "ms/step", round(d["ms_per_step"],3), "configs1", (d.get("configs1") or {}).get("ms_per_step"))
    except Exception as e:
        print(n, "no line", e)
PY
