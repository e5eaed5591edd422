"""Data-parallel training step on real GPUs vs the single-GPU step on the concatenated batch.

  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/train_dp_check.py

Every rank first runs the full-batch step alone (no process group), then the ranks run the sharded step with the NCCL
gradient / BatchNorm-statistics exchange; losses, updated weights and BatchNorm running statistics must agree.
DP_GRAPH=1: the sharded side runs through graph_step.CapturedTrainStep (one eager warm-up step + one replay of the captured
graph, NCCL collectives inside the graph) and the reference side takes the same two steps eagerly."""
import copy, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from dex_gan_grasp_b200 import dataparallel, synth
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
import bench

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
B = int(os.environ.get("DP_BATCH", "512"))
prec = os.environ.get("DEXGG_PRECISION", "fp32")
graph = os.environ.get("DP_GRAPH", "0") == "1"
combined = os.environ.get("DP_COMBINED", "0") == "1"  # GAN + evaluator update per step (TrainEngine.combined_step: two chains)
wgan = os.environ.get("DP_WGAN", "0") == "1"          # WGAN-GP variant (is_wgan): the gradient-penalty pass with cross-rank BatchNorm statistics


def fresh():
    m = DexGanGrasp(dict(bench.TRAIN_CFG, gpu_ids=[local], precision=prec, is_wgan=wgan))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(7))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(7))
    if combined:
        m.DexEvaluator.dout.p = 0.0  # the dropout stream is per local row: switched off so that sharded == full batch
    return m


full = synth.synth_train_batch(B, seed=11)
z = synth.synth_latents(B, 5, seed=3)
alphas = None
if wgan:  # the three interpolation weights of the penalty (the reference draws them per call): fixed here, sharded like the batch
    ga = torch.Generator().manual_seed(17)
    alphas = [torch.randn((B,) + (1,) * (full[k].dim() - 1), generator=ga) for k in ("rot_matrix", "transl", "joint_conf")]
ref = fresh()
for _ in range(2 if graph else 1):
    lg = ref.update_dexgangrasp({k: v.to(dev) for k, v in full.items()}, is_train_gen=True, Zgen=z.to(dev), alphas=alphas)
    if combined:
        lg = dict(lg, **ref.update_dexevaluator({k: v.to(dev) for k, v in full.items()}))
torch.cuda.synchronize()

dist.init_process_group("nccl", device_id=dev)
dp = fresh()
n = B // world
mine = {k: v.to(dev) for k, v in dataparallel.shard(full, rank, world).items()}
zmine = z[rank * n:(rank + 1) * n].to(dev)
cap = None
if graph:
    from dex_gan_grasp_b200.graph_step import CapturedTrainStep
    cap = CapturedTrainStep(dp, mine, with_evaluator=combined, is_train_gen=True, warmup=1, example_z=zmine)
    lg2 = cap(mine, z=zmine)
else:
    lg2 = dp.update_dexgangrasp(mine, is_train_gen=True, Zgen=zmine, alphas=None if alphas is None else [a[rank * n:(rank + 1) * n] for a in alphas])
    if combined:
        lg2 = dict(lg2, **dp.update_dexevaluator(mine))
torch.cuda.synchronize()
out = {"world": world, "batch": B, "precision": prec, "captured_graph": graph, "combined": combined, "wgan": wgan,
       "engine": dp.train_engine_name(), "collectives": dataparallel.collective_counters(), "loss_abs_diff": {}, "max_weight_diff": 0.0, "max_buffer_diff": 0.0}
for k in lg:
    # per-rank losses are means over the local shard: average them over ranks to compare with the full-batch mean
    t = torch.tensor([float(lg2[k])], device=dev, dtype=torch.float64)
    dist.all_reduce(t)
    out["loss_abs_diff"][k] = abs(float(t) / world - float(lg[k]))
params = {k for k, _ in ref.DexGANGrasp.named_parameters()}
sd_a, sd_b = ref.DexGANGrasp.state_dict(), dp.DexGANGrasp.state_dict()
if combined:
    params |= {"eva." + k for k, _ in ref.DexEvaluator.named_parameters()}
    sd_a = dict(sd_a, **{"eva." + k: v for k, v in ref.DexEvaluator.state_dict().items()})
    sd_b = dict(sd_b, **{"eva." + k: v for k, v in dp.DexEvaluator.state_dict().items()})
for k in sd_a:
    if not sd_a[k].is_floating_point():
        continue
    d = float((sd_a[k].double() - sd_b[k].double()).abs().max())
    key = "max_weight_diff" if k in params else "max_buffer_diff"
    out[key] = max(out[key], d)
# gradients left by the step: the averaged sharded gradients must equal the full-batch gradients (relative L2 norm, worst tensor)
ga, gb = dict(ref.DexGANGrasp.named_parameters()), dict(dp.DexGANGrasp.named_parameters())
worst = (0.0, "")
for k in ga:
    if ga[k].grad is None or gb[k].grad is None or k.endswith(("fc1.bias", "fc2.bias")):
        continue
    den = float(ga[k].grad.double().norm())
    if den > 1e-9:
        worst = max(worst, (float((ga[k].grad.double() - gb[k].grad.double()).norm()) / den, k))
out["worst_grad_rel_diff"] = worst
# weights must be identical on every rank after the step
flat = torch.cat([p.detach().reshape(-1) for p in list(dp.DexGANGrasp.parameters()) + list(dp.DexEvaluator.parameters())])
mx, mn = flat.clone(), flat.clone()
dist.all_reduce(mx, op=dist.ReduceOp.MAX)
dist.all_reduce(mn, op=dist.ReduceOp.MIN)
out["max_rank_divergence"] = float((mx - mn).abs().max())
if rank == 0:
    print(json.dumps(out), flush=True)
if cap is not None:
    cap.close()
import threading, time
threading.Thread(target=lambda: (time.sleep(20), os._exit(0)), daemon=True).start()
dist.destroy_process_group()
