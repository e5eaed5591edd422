#!/bin/bash
# One node, N GPUs: data-parallel parity check, the C3 bench line, the C5 training step (strong and weak).
# usage: bash tools/dp_scale.sh N TAG [nobench]  (writes gpurun_out/TAG_*_nN.json)
N=$1; TAG=$2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
DP_GRAPH=1 DP_COMBINED=1 DP_BATCH=4096 DEXGG_PRECISION=bf16 timeout 300 $TR tools/train_dp_check.py 2>/dev/null | grep '^{' | tail -1 > gpurun_out/${TAG}_dp_check_n$N.json
[ "$3" = nobench ] || timeout 600 $TR bench.py --gpus $N 2>/dev/null | tail -1 > gpurun_out/${TAG}_bench_n$N.json
timeout 300 $TR bench.py --gpus $N --workload train --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/${TAG}_train_strong_n$N.json
timeout 300 $TR bench.py --gpus $N --workload train --train-scaling weak --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/${TAG}_train_weak_n$N.json
python - <<PY
import json, os
for f in ["dp_check", "bench", "train_strong", "train_weak"]:
    p = "gpurun_out/${TAG}_%s_n$N.json" % f
    if not os.path.exists(p):
        continue
    try:
        d = json.loads(open(p).read())
        print(f, {k: d.get(k) for k in ("value", "ms_per_step", "max_rank_divergence", "max_weight_diff") if k in d},
              (d.get("strong") or {}).get("ms") if f == "bench" else "")
    except Exception as e:
        print(f, "ERR", e, open(p).read()[-300:])
PY
