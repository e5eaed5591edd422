"""Kernel timeline of the training step (torch.profiler / CUPTI; there is no nsys on these boxes): per-kernel totals, NCCL time
and GPU idle gaps of the eager and the captured step.  torchrun --nproc-per-node N tools/train_timeline.py [--graph]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile
from dex_gan_grasp_b200 import dataparallel, synth
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
import bench

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
B = int(os.environ.get("TL_BATCH", "4096"))
graph = "--graph" in sys.argv
m = DexGanGrasp(dict(bench.TRAIN_CFG, gpu_ids=[local], precision="bf16"))
m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(7))
m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(7))
full = synth.synth_train_batch(B, seed=11)
mine = {k: v.to(dev) for k, v in dataparallel.shard(full, rank, world).items()}
rows = B // world
eng = m._engine(rows)
z = torch.randn((rows, 5), device=dev)
if graph:
    from dex_gan_grasp_b200.graph_step import CapturedTrainStep
    cap = CapturedTrainStep(m, mine, warmup=2)
    step = lambda: cap(mine)
else:
    step = lambda: eng.combined_step(dict(mine), z, True)
for _ in range(3):
    step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
if rank == 0 and evs:
    t0, t1 = evs[0].time_range.start, max(e.time_range.end for e in evs)
    busy, last_end, gaps = 0.0, t0, 0.0
    agg = {}
    for e in evs:
        d = e.time_range.end - e.time_range.start
        a = agg.setdefault(e.name[:60], [0, 0.0])
        a[0] += 1
        a[1] += d
        if e.time_range.start > last_end:
            gaps += e.time_range.start - last_end
        last_end = max(last_end, e.time_range.end)
    print(json.dumps({"world": world, "rows_per_gpu": rows, "graph": graph, "steps": 3, "span_ms_per_step": (t1 - t0) / 3e3, "idle_ms_per_step": gaps / 3e3,
                      "kernels_per_step": len(evs) / 3}))
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:28]:
        print(f"{t / 3e3:8.3f} ms/step  {n / 3:6.1f} x  {k}")
if world > 1:
    import threading, time
    threading.Thread(target=lambda: (time.sleep(15), os._exit(0)), daemon=True).start()
    dist.destroy_process_group()
