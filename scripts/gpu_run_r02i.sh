#!/bin/bash
# Round-2 ninth GPU pass: seq_attention at 4 CTAs/SM (tests, timings, bench), full parity suite.
O=gpurun_out/r02i; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -q ) > $O/pytest_all.log 2>&1; echo "rc=$?" >> $O/pytest_all.log
tail -4 $O/pytest_all.log
timeout 300 python scripts/gpu_ab_act.py default > $O/ab_act.jsonl 2> $O/ab_act.err; grep attention $O/ab_act.jsonl
( timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.log").read().strip().splitlines()[-1])
a=d["acceptance"]["agree_by_construction"]
print("ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "agree", a["embed_shift"], a["draft_acceptance_rate"], round(a["value"],1), "configs1", (d.get("configs1") or {}).get("ms_per_step"))
PY
timeout 300 python scripts/gpu_timeline.py --seqs 64 > $O/timeline_64.md 2> $O/timeline_64.err
timeout 300 python scripts/gpu_timeline.py --seqs 4 > $O/timeline_4.md 2> $O/timeline_4.err
ls -la $O
