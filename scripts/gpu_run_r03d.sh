#!/bin/bash
# Round-2 final pass (1 GPU): full GPU suite, smoke(), default bench line, reference arm.
O=gpurun_out/r03d; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -6 $O/pytest_all.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" ) > $O/smoke.log 2>&1; tail -2 $O/smoke.log
( timeout 900 python bench.py ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.log").read().strip().splitlines()[-1])
a=d["acceptance"]["agree_by_construction"]
print("ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "agree", a["embed_shift"], a["draft_acceptance_rate"], round(a["value"],1), "configs1", d["configs1"]["ms_per_step"], "steps", d["steps"], "warmup", d["warmup"])
print(d["roofline"]["frac"], d["roofline"]["traffic"], d["clocks"], d["gpu_launches"])
for w in ("workload","configs1"):
    print({k:(v["w4a4"]["us"], v["w4a16"]["us"]) for k,v in d["per_layer"][w].items()})
PY
