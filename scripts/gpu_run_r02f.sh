#!/bin/bash
# Round-2 sixth GPU pass: seq_attention with staged values, down-rotation cluster form for many rows, attention-path A/B in the bench.
O=gpurun_out/r02f; mkdir -p $O
( timeout 300 python -m pytest tests/test_gpu_model.py -m gpu -q -k "seq_attention or attn_block" ) > $O/pytest_attn.log 2>&1; echo "rc=$?" >> $O/pytest_attn.log
tail -3 $O/pytest_attn.log
timeout 300 python scripts/gpu_ab_act.py default > $O/ab_act.jsonl 2> $O/ab_act.err
QSPEC_K28_CLUSTER_ROWS=4096 timeout 300 python scripts/gpu_ab_act.py k28cluster >> $O/ab_act.jsonl 2>> $O/ab_act.err
cat $O/ab_act.jsonl
( timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs1 --agree-shifts 18 ) > $O/bench_default.log 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
( QSPEC_ATTN_BLOCK_MAX_T=32 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs1 --no-per-layer --agree-shifts 20 ) > $O/bench_seqattn_draft.log 2> $O/bench_seqattn_draft.err; echo "rc=$?" >> $O/bench_seqattn_draft.err
( QSPEC_K28_CLUSTER_ROWS=4096 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs1 --no-per-layer --agree-shifts 22 ) > $O/bench_k28cluster.log 2> $O/bench_k28cluster.err; echo "rc=$?" >> $O/bench_k28cluster.err
for f in default seqattn_draft k28cluster; do python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$f.log").read().strip().splitlines()[-1])
    a=d["acceptance"]["agree_by_construction"]
    print("$f", "ms/step", round(d["ms_per_step"],3), "agree", a["embed_shift"], a["draft_acceptance_rate"], round(a["value"],1))
except Exception as e: print("$f", "ERR", e)
PY
done
ls -la $O
