#!/bin/bash
# C3 (weak scaling) and C5 (strong scaling) at N = 1, 2, 4, 8 on one box; one JSON line per run into gpurun_out/scale_*.jsonl
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > gpurun_out/scale_c3.jsonl; : > gpurun_out/scale_train_fp32.jsonl; : > gpurun_out/scale_train_bf16.jsonl
for n in ${NS:-1 2 4 8}; do
  timeout -k 5 240 $T --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --no-fp32-line --no-cpu-baseline 2>/dev/null | tail -1 >> gpurun_out/scale_c3.jsonl
  timeout -k 5 240 $T --nproc-per-node $n --master-port $((29700+n)) bench.py --gpus $n --workload train --precision fp32 --steps 10 --warmup 3 2>/dev/null | tail -1 >> gpurun_out/scale_train_fp32.jsonl
  timeout -k 5 240 $T --nproc-per-node $n --master-port $((29800+n)) bench.py --gpus $n --workload train --precision bf16 --steps 10 --warmup 3 2>/dev/null | tail -1 >> gpurun_out/scale_train_bf16.jsonl
done
python - <<'PY'
import json
for f in ("c3", "train_fp32", "train_bf16"):
    for l in open(f"gpurun_out/scale_{f}.jsonl"):
        try:
            d = json.loads(l); print(f, d["n_gpus"], round(d["value"]), d["unit"], round(d["ms_per_step"], 3), "ms", "e2e", round(d["e2e"]["value"]))
        except Exception as e:
            print(f, "bad line", l[:100])
PY
