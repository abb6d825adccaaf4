#!/bin/bash
# Round-2 pass 32 (1 GPU): bench A/B of two scheduler knobs on the final code (64 sequences): k28 one-CTA-per-row form above 64 rows;
# CTA pairs from 64 tokens (the int8-container draft GEMMs).
O=gpurun_out/r03f; mkdir -p $O
run() {
  ( env "$@" timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-per-layer --no-configs1 --agree-shifts 19 ) > $O/bench_$TAG.log 2> $O/bench_$TAG.err
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$TAG.log").read().strip().splitlines()[-1])
    print("$TAG ms/step", round(d["ms_per_step"],3), "acceptance", d["acceptance"]["agree_by_construction"]["draft_acceptance_rate"])
except Exception as e:
    print("$TAG no line", e)
PY
}
TAG=default run QSPEC_NOOP=1
TAG=k28_plain_above_64 run QSPEC_K28_CLUSTER_ROWS=64
TAG=pairs_from_64 run QSPEC_GEMM_PAIR_MIN_M=64
TAG=default2 run QSPEC_NOOP=1
