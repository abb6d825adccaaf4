#!/bin/bash
# Round-2 eighth GPU pass: tensor-core seq_attention (tests, timings), bench A/B of the attention path, ncu of the dominant GEMMs.
O=gpurun_out/r02h; mkdir -p $O
( timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_tp_onegpu.py tests/test_gpu_qwen.py -m gpu -q ) > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
tail -4 $O/pytest.log
timeout 300 python scripts/gpu_ab_act.py default > $O/ab_act.jsonl 2> $O/ab_act.err; cat $O/ab_act.jsonl
( timeout 900 python bench.py --steps 20 --warmup 3 ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
( QSPEC_ATTN_BLOCK_MAX_T=1024 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs1 --no-per-layer --agree-shifts 19 ) > $O/bench_attnblock_draft.log 2> $O/bench_attnblock_draft.err; echo "rc=$?" >> $O/bench_attnblock_draft.err
for f in default attnblock_draft; do python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$f.log").read().strip().splitlines()[-1])
    a=d["acceptance"]["agree_by_construction"]
    print("$f", "ms/step", round(d["ms_per_step"],3), "agree", a["embed_shift"], a["draft_acceptance_rate"], round(a["value"],1), "configs1", (d.get("configs1") or {}).get("ms_per_step"))
except Exception as e: print("$f", "ERR", e)
PY
done
timeout 300 python scripts/gpu_timeline.py --seqs 64 > $O/timeline_64.md 2> $O/timeline_64.err
CMD="python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline --no-configs1 --no-per-layer --agree-shifts="
$CMD > $O/plain2.log 2>&1 &&
ncu --nvtx --nvtx-include "timed/" --set full --clock-control none --import-source on -k regex:w4_stream -s 386 -c 1 -o $O/prof_w4a16_m256_gate_up $CMD > $O/ncu_w4a16.log 2>&1
ncu --nvtx --nvtx-include "timed/" --set full --clock-control none --import-source on -k regex:w4_stream -s 2 -c 1 -o $O/prof_w4a4_m64_gate_up $CMD > $O/ncu_w4a4.log 2>&1
ncu --nvtx --nvtx-include "timed/" --set full --clock-control none --import-source on -k regex:seq_attention -s 2 -c 1 -o $O/prof_seq_attention $CMD > $O/ncu_seqattn.log 2>&1
ls -la $O
