set -x
cd /root/repo
export CUDA_DEVICE_MAX_CONNECTIONS=32
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
DP_BATCH=4096 DEXGG_PRECISION=bf16 timeout 300 $TR tools/train_dp_check.py 2>&1 | grep -v Warning | tail -3
DP_GRAPH=1 DP_COMBINED=1 DP_BATCH=4096 DEXGG_PRECISION=bf16 timeout 300 $TR tools/train_dp_check.py 2>&1 | grep -v Warning | tail -3
DP_COMBINED=1 DP_BATCH=4096 DEXGG_PRECISION=bf16 timeout 300 $TR tools/train_dp_check.py 2>&1 | grep -v Warning | tail -3
timeout 300 $TR bench.py --gpus 2 --workload train --steps 30 --warmup 5 2>&1 | tail -1 > gpurun_out/r02l_train_strong_n2.json
timeout 300 $TR bench.py --gpus 2 --workload train --train-scaling weak --steps 30 --warmup 5 2>&1 | tail -1 > gpurun_out/r02l_train_weak_n2.json
DEXGG_DP_GRADS=nccl timeout 300 $TR bench.py --gpus 2 --workload train --train-scaling weak --steps 30 --warmup 5 2>&1 | tail -1 > gpurun_out/r02l_train_weak_n2_ncclgrads.json
python - <<'PY'
import json
for f in ["r02l_train_strong_n2","r02l_train_weak_n2","r02l_train_weak_n2_ncclgrads"]:
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read())
        print(f, d["ms_per_step"], d["value"], d.get("collectives_per_step"))
    except Exception as e:
        print(f, "ERR", e, open("gpurun_out/%s.json"%f).read()[-500:])
PY
