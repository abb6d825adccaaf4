#!/bin/bash
# Round-2 tenth GPU pass: 256-token tiles (one weight pass at M = 256) -- parity, A/B, bench.
O=gpurun_out/r02j; mkdir -p $O
L=$PWD/qspec_b200/_C
( timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_full_sizes.py -m gpu -q -x -k "w4a16" ) > $O/pytest_w4a16.log 2>&1; rc=$?; echo "rc=$rc" >> $O/pytest_w4a16.log
tail -4 $O/pytest_w4a16.log
if [ $rc -ne 0 ]; then ls -la $O; exit 0; fi
QSPEC_GEMM_NP256=0 timeout 300 python scripts/gpu_ab.py np128 > $O/ab.jsonl 2> $O/ab.err
QSPEC_GEMM_NP256=1 timeout 300 python scripts/gpu_ab.py np256 >> $O/ab.jsonl 2>> $O/ab.err
grep w4a16 $O/ab.jsonl
( time timeout 900 python -m pytest tests -m gpu -q ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -4 $O/pytest_all.log
( timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.log").read().strip().splitlines()[-1])
a=d["acceptance"]["agree_by_construction"]
print("ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1), "agree", a["embed_shift"], a["draft_acceptance_rate"], round(a["value"],1), "configs1", (d.get("configs1") or {}).get("ms_per_step"))
print({k:(v["w4a4"]["us"], v["w4a16"]["us"]) for k,v in d["per_layer"]["workload"].items()})
PY
for spec in "w4a16 256 28672 4096"; do
  echo "==== $spec" >> $O/gemm_trace.txt
  QSPEC_LIB=$L/libqspec_b200_prof.so timeout 120 python scripts/gpu_gemm_trace.py $spec >> $O/gemm_trace.txt 2>&1
done
ls -la $O
