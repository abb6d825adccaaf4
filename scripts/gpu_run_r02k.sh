#!/bin/bash
# Round-2 2-GPU pass: TP=2 test (fused reduce vs NCCL vs TP oracle), configs[2] at TP=2 (64 sequences in total), reference arm at N=2.
O=gpurun_out/r02k; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
( time timeout 600 python -m pytest tests/test_gpu_tp.py tests/test_gpu_kernels.py -m gpu -q -x -k "tp or w4a16_exact" ) > $O/pytest_tp.log 2>&1; echo "rc=$?" >> $O/pytest_tp.log
tail -15 $O/pytest_tp.log
( timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 3 ) > $O/bench_n2.log 2> $O/bench_n2.err; echo "rc=$?" >> $O/bench_n2.err
tail -3 $O/bench_n2.err
tail -c 3000 $O/bench_n2.log
( timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline ) > $O/bench_n1.log 2> $O/bench_n1.err; echo "rc=$?" >> $O/bench_n1.err
python - <<PY
import json
for n in (1, 2):
    try:
        d=json.loads(open("$O/bench_n%d.log" % n).read().strip().splitlines()[-1])
        print(n, "ms/step", round(d["ms_per_step"],3), "value", round(d["value"],1), "e2e", d["e2e"]["value"], "agree", (d.get("acceptance") or {}).get("agree_by_construction"))
    except Exception as e:
        print(n, "no line", e)
PY
ls -la $O
