#!/bin/bash
# Round-2 8-GPU pass: BASELINE configs[2] at TP = 8 and 4 (64 sequences in total), configs[3] (70B, TP = 8, 32 sequences).
# One bench at a time; each under its own timeout.
O=gpurun_out/${QSPEC_RUN_TAG:-r02n}; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
run() {  # name, nproc, extra args
  local name=$1 n=$2; shift 2
  ( timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $n --steps 20 --warmup 3 "$@" ) > $O/bench_$name.log 2> $O/bench_$name.err
  echo "rc=$?" >> $O/bench_$name.err
  tail -2 $O/bench_$name.err
}
run n8 8
run n4 4
run 70b_n8 8 --model 70b --seqs 32
python - <<PY
import json
for n in ("n8", "n4", "70b_n8"):
    try:
        d=json.loads(open("$O/bench_%s.log" % n).read().strip().splitlines()[-1])
        a=(d.get("acceptance") or {}).get("agree_by_construction") or {}
        print(n, "ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "agree", a.get("embed_shift"), a.get("draft_acceptance_rate"), a.get("value"))
        print("  ", {k:(v["w4a4"]["us"], v["w4a16"]["us"]) for k,v in d["per_layer"]["workload"].items()})
    except Exception as e:
        print(n, "no line", e)
PY
ls -la $O
