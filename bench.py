#!/usr/bin/env python
"""bench.py — grasps/sec of the DexGANGrasp hot path (BPS -> DexGenerator -> DexEvaluator -> top-k) on B200.

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): 1024 synthetic single-view point clouds of
2048 points x 1000 latent samples per cloud -> evaluator -> per-object top-10 above threshold 0.5.  One process per GPU.

  headline `value`   every rank runs 1024 clouds of its own (no data-path collective) -> "scaling": "weak"
  `strong`           the config read literally: 1024 clouds IN TOTAL sharded by object over the N ranks
                     (dataparallel.shard_objects; z of object i from dataparallel.object_seed, independent of N)
  `train`            C5 (configs[4]): GAN + evaluator training step, global batch 4096, bf16, data parallel over NCCL
  `parity_check`     the first 16 objects of the timed workload against the fp32 oracle pipeline (outside every timed region)
  `gpu_baseline`     the reference algorithm as torch-eager fp32 ON THE GPU (TF32 off): the "existing GPU path" 
  `cpu_baseline`     the reference algorithm on the box's host cores (oracle/: C BPS + torch-CPU networks)

  python bench.py [--gpus N] [--steps K] [--warmup W] [--precision fp32|bf16] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...

`--impl reference` times the CPU restatement of the reference path on a bounded sample of the same workload.
Prints ONE JSON line on rank 0.  Per-step times are CUDA-event times on the launching stream; `ms_per_step` is their MEDIAN
(p90 / mean / min / max beside it), max over ranks.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

N_CLOUDS, N_POINTS, N_SAMPLES, TOPK, THRESH, SEED, Z_SEED = 1024, 2048, 1000, 10, 0.5, 7, 1234
MAC_PER_GRASP, MAC_PER_OBJECT = 1_881_600, 15_728_640  # hoisted algorithm (SURVEY.md §8a note H)
# C5 algorithmic FLOP per row: GAN step 153.5 MFLOP (fwd G + 3 D, D-step bwd, G-step bwd), evaluator 63.9
TRAIN_FLOP_PER_ROW_GAN, TRAIN_FLOP_PER_ROW_EVA = 153.5e6, 63.9e6
METRIC = "grasps/sec (BPS->DexGenerator->DexEvaluator->top-k)"
WORKLOAD = "C3 batched inference: synthetic clouds (2048 pts) x 1000 latent samples -> evaluator top-10"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("DEXGG_PRECISION", "bf16"), choices=["fp32", "bf16"],
                    help="MLP arithmetic: bf16 = tcgen05 tensor cores (headline, north_star), fp32 = CUDA-core 1e-4 mode")
    ap.add_argument("--no-fp32-line", action="store_true", help="skip the secondary fp32-exact-mode measurement")
    ap.add_argument("--clouds", type=int, default=N_CLOUDS)
    ap.add_argument("--samples", type=int, default=N_SAMPLES)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the C5 training block of the default line")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-c1", action="store_true", help="skip the one-object latency block")
    ap.add_argument("--workload", default="c3", choices=["c3", "train"],
                    help="c3 = the headline inference metric (with the C5 block inside); train = C5 as its own line")
    ap.add_argument("--batch", type=int, default=4096, help="global training batch (C5)")
    ap.add_argument("--train-scaling", default="strong", choices=["strong", "weak"],
                    help="C5: strong = the global batch is split over the GPUs (SURVEY 8e); weak = --batch rows on every GPU")
    ap.add_argument("--train-steps", type=int, default=30)
    ap.add_argument("--no-graph", action="store_true", help="C5: issue the step eagerly instead of replaying the captured CUDA graph")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json"))), "MEASURED_PEAKS.json"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


def stats(ms_list):
    a = np.asarray(ms_list, dtype=np.float64)
    return {"median": float(np.median(a)), "p90": float(np.percentile(a, 90)), "mean": float(a.mean()), "min": float(a.min()),
            "max": float(a.max()), "n": int(a.size)}


def object_latents(obj_ids, n_samples, zD=5):
    """z of every object from seed (+) object id (dataparallel.object_seed): the same object gets the same latents whatever the
    number of GPUs."""
    from dex_gan_grasp_b200 import dataparallel
    out = torch.empty((len(obj_ids) * n_samples, zD), dtype=torch.float32)
    g = torch.Generator()
    for j, o in enumerate(obj_ids):
        g.manual_seed(dataparallel.object_seed(Z_SEED, o))
        out[j * n_samples:(j + 1) * n_samples] = torch.randn((n_samples, zD), generator=g)
    return out


def object_clouds(obj_ids, n_points):
    """Cloud of global object id i (seed 1000 + i; ids beyond the first 64 re-pose one of those shapes)."""
    from dex_gan_grasp_b200 import synth
    return synth.synth_clouds_by_id(obj_ids, n_points, seed0=1000)


# --------------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference_sample(n_objects, n_samples, n_points, obj0=0, repeats=1):
    """The reference's per-object loop (dexgangrasp_offline.py:147-162 / dexgangrasp_real_robot.py:139-150) restated on
    the host: BPS (oracle/bps_oracle.c, all cores) -> np.tile -> generator -> rot6d -> palm offset -> evaluator ->
    sort / threshold.  Returns (grasps, seconds, cores)."""
    from dex_gan_grasp_b200 import synth
    from oracle import nets_oracle as no
    so = os.path.join(REPO, "oracle", "libdexgg_oracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", os.path.join(REPO, "oracle")])
    lib = ctypes.CDLL(so)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    basis = np.load(os.path.join(REPO, "dex_gan_grasp_b200", "basis_point_set.npy")).astype(np.float32)
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    ids = list(range(obj0, obj0 + n_objects))
    clouds = list(synth.synth_clouds_by_id(ids, n_points))
    zs = object_latents(ids, n_samples)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        kept = 0
        for j, c in enumerate(clouds):
            d = np.empty(4096, np.float32)
            idx = np.empty(4096, np.int32)
            lib.dexgg_oracle_bps_encode_mt(vp(basis), 4096, vp(c), len(c), vp(d), vp(idx), min(cores, 256))
            bps_t = torch.from_numpy(np.tile(d, (n_samples, 1)))
            g = no.generate_poses(sd_gan, zs[j * n_samples:(j + 1) * n_samples], bps_t, 0.025)
            try:
                f, _ = no.filter_grasps(sd_eva, d, g, THRESH)
                kept += min(TOPK, len(f["transl"]))
            except ValueError:
                pass
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_objects * n_samples, best, cores


def run_reference(args, rank):
    if rank != 0:
        return
    n_obj = 16  # "C3 on the first >= 16 objects x 1000 z"
    times = []
    for i in range(args.warmup + args.steps):
        g, dt, cores = cpu_reference_sample(n_obj, args.samples, N_POINTS, obj0=(16 * i) % 1024)
        if i >= args.warmup:
            times.append(dt * 1e3)
    st = stats(times)
    ms = st["median"]
    val = n_obj * args.samples / (ms * 1e-3)
    sample = f"{n_obj} objects x {args.samples} z per step (of {args.clouds} x {args.samples}), oracle/ C BPS + torch-CPU networks"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "grasps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "ms_per_step_stats": st, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "clouds_per_gpu": args.clouds, "points": N_POINTS, "samples_per_cloud": args.samples,
                       "topk": TOPK, "thresh": THRESH},
            "cpu_baseline": {"value": val, "unit": "grasps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "grasps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- torch-eager GPU baseline
def gpu_reference_sample(n_objects, n_samples, n_points, dev, repeats=3):
    """The reference ALGORITHM as it would run on this GPU with stock torch (SURVEY.md §8d "existing GPU path"): per object,
    bps_torch-style encode (broadcast distances -> argmin -> gather -> sqrt, torch ops), np.tile of the encoding + H2D copy
    (dexgangrasp.py:509-512), Generator.forward as ~25 ATen kernels, rot6d, D2H to numpy + palm offset (networks.py:246-253),
    then filter_grasps: tile again, H2D, DexEvaluator.forward, sort, gather, D2H (dexgangrasp.py:393-442).  fp32, TF32 off.
    oracle/nets_oracle.py supplies the functional restatement of the nn.Modules (it is device-agnostic).  Baseline leg only."""
    from dex_gan_grasp_b200 import synth
    from oracle import nets_oracle as no
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd_gan = {k: v.to(dev) for k, v in synth.synth_gan_state_dict(SEED).items()}
    sd_eva = {k: v.to(dev) for k, v in synth.synth_eva_state_dict(SEED).items()}
    basis = torch.from_numpy(np.load(os.path.join(REPO, "dex_gan_grasp_b200", "basis_point_set.npy")).astype(np.float32)).to(dev)
    ids = list(range(n_objects))
    clouds = list(synth.synth_clouds_by_id(ids, n_points))
    zs = object_latents(ids, n_samples)
    best = None
    for _ in range(repeats + 1):
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        kept = 0
        with torch.no_grad():
            for j, c in enumerate(clouds):
                x = torch.from_numpy(c).to(dev)                                   # bps.encode(pc_tensor)
                d2 = ((basis[:, None, :] - x[None, :, :]) ** 2).sum(-1)
                nn_idx = d2.argmin(1)
                enc = torch.sqrt(((x[nn_idx] - basis) ** 2).sum(1)).cpu().numpy()  # enc_dict['dists'].cpu().numpy()
                bps_t = torch.tensor(np.tile(enc, (n_samples, 1)), device=dev)     # np.tile + H2D (dexgangrasp.py:509-512)
                z = zs[j * n_samples:(j + 1) * n_samples].to(dev)
                out = no.generator_forward(sd_gan, z, bps_t)
                rot = no.rot_matrix_from_ortho6d(out["rot_6D"])
                g = {"rot_matrix": rot.cpu().numpy(), "transl": out["transl"].cpu().numpy(), "joint_conf": out["joint_conf"].cpu().numpy()}
                g["transl"] = no.translate_along_axis(g["transl"], g["rot_matrix"], 0, 0.025)
                bps_t2 = torch.tensor(np.tile(enc, (n_samples, 1)), device=dev)    # filter_grasps tiles and copies again (:403-410)
                p = no.evaluator_forward(sd_eva, bps_t2, torch.tensor(g["rot_matrix"], device=dev), torch.tensor(g["transl"], device=dev)).squeeze(-1)
                sp, si = p.sort(descending=True)
                keep = si[sp > THRESH]
                res = {k: torch.tensor(v, device=dev)[keep].cpu().numpy() for k, v in g.items()}
                kept += min(TOPK, len(res["transl"]))
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_objects * n_samples, best


# --------------------------------------------------------------------------------------------- C1: one object, online latency
def c1_latency(model, dev, n_iter=200):
    """Config C1 / the online shape of dexgangrasp_real_robot.py:131-162: ONE real segmented cloud (soft_scrub, 38 663 points,
    tests/golden/real_clouds_full.npz = the reference's data/real_objects/object/soft_scrub.pcd) -> BPS -> S grasps -> evaluator
    -> kept grasps, host buffers in and out.  Two ways through the public API:
      class_surface  BPSEncoder.encode_pcd_with_bps -> DexGanGrasp.generate_grasps(S) -> filter_grasps(thresh) with numpy in /
                     out exactly like the robot script (four host round trips)
      pipeline       GraspPipeline.run_host (B = 1, one captured CUDA graph, one H2D of the cloud + z, one D2H of the top-10)
    p50 / p99 wall-clock per call, against the CPU port of the same sequence."""
    from dex_gan_grasp_b200 import bps as bps_mod
    from dex_gan_grasp_b200.pipeline import GraspPipeline
    path = os.path.join(REPO, "tests", "golden", "real_clouds_full.npz")
    if not os.path.exists(path):
        return {"error": "tests/golden/real_clouds_full.npz missing"}
    cloud = np.load(path)["soft_scrub/cloud"]
    enc = bps_mod.BPSEncoder({"gpu_ids": [dev.index]})
    out = {"cloud": "soft_scrub.pcd (reference data/real_objects), %d points" % len(cloud), "iterations": n_iter}
    pipe = GraspPipeline(model.DexGANGrasp, model.DexEvaluator, k=TOPK, thresh=THRESH, z_offset=0.025, use_graph=True)
    cloud_h = torch.from_numpy(cloud).unsqueeze(0).pin_memory()
    for S in (400, 1000):
        rec = {}
        # -- class surface, numpy in / out
        ts = []
        pts64 = cloud.astype(np.float64)
        for i in range(n_iter // 4 + 3):
            t0 = time.perf_counter()
            b = enc.encode_pcd_with_bps(pts64)
            g = model.generate_grasps(b, n_samples=S, return_arr=True)
            try:
                f = model.filter_grasps(b, g, thresh=THRESH)
                kept = len(f["transl"])
            except ValueError:
                kept = 0
            ts.append(time.perf_counter() - t0)
        ts = np.array(ts[3:]) * 1e3
        rec["class_surface_ms"] = {"p50": float(np.percentile(ts, 50)), "p99": float(np.percentile(ts, 99)), "kept": kept}
        # -- batched pipeline API with B = 1 (captured graph)
        z_h = torch.randn((S, 5), generator=torch.Generator().manual_seed(5)).pin_memory()
        for _ in range(5):
            pipe.run_host(cloud_h, z_h, S)
        ts = []
        for i in range(n_iter):
            t0 = time.perf_counter()
            res, h2d, d2h = pipe.run_host(cloud_h, z_h, S)
            ts.append(time.perf_counter() - t0)
        ts = np.array(ts) * 1e3
        rec["pipeline_ms"] = {"p50": float(np.percentile(ts, 50)), "p99": float(np.percentile(ts, 99)), "kept": int(res["count"].sum()),
                              "h2d_bytes": h2d, "d2h_bytes": d2h}
        out["S%d" % S] = rec
    return out


def c1_cpu_port(S=400):
    """The same one-object sequence on the host cores (oracle/: C BPS + torch-CPU networks)."""
    from dex_gan_grasp_b200 import synth
    from oracle import nets_oracle as no
    lib = ctypes.CDLL(os.path.join(REPO, "oracle", "libdexgg_oracle.so"))
    cloud = np.load(os.path.join(REPO, "tests", "golden", "real_clouds_full.npz"))["soft_scrub/cloud"]
    basis = np.load(os.path.join(REPO, "dex_gan_grasp_b200", "basis_point_set.npy")).astype(np.float32)
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ts = []
    for i in range(6):
        t0 = time.perf_counter()
        d, idx = np.empty(4096, np.float32), np.empty(4096, np.int32)
        lib.dexgg_oracle_bps_encode_mt(vp(basis), 4096, vp(cloud), len(cloud), vp(d), vp(idx), min(cores, 256))
        z = torch.randn((S, 5))
        g = no.generate_poses(sd_gan, z, torch.from_numpy(np.tile(d, (S, 1))), 0.025)
        try:
            no.filter_grasps(sd_eva, d, g, THRESH)
        except ValueError:
            pass
        ts.append(time.perf_counter() - t0)
    return {"S": S, "ms_p50": float(np.median(ts[1:]) * 1e3), "cores": cores}


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- C5: training step
TRAIN_CFG = dict(latentD=5, n_hand_joints=12, is_train=True, model="dexgangrasp", is_wgan=False, bce_weight=1.0,
                 train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
                 lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
                 init_gain=0.02)


def train_measure(args, rank, world, local, precision, steps, warmup, scaling="strong", graph=True):
    """C5 (BASELINE.json configs[4]): one step = DexGanGrasp.update_dexgangrasp (D step + G step) + update_dexevaluator on a
    global batch of 4096 synthetic rows, sharded 4096/N per rank, gradients and BatchNorm statistics exchanged over NCCL
    (train.py:233-260 drives the same two calls per batch).  Returns the `train` record (rank 0) or None."""
    import torch.distributed as dist
    from dex_gan_grasp_b200 import _native, dataparallel, synth, train_ops
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    L = _native.lib()
    dev = torch.device("cuda", local)
    model = DexGanGrasp(dict(TRAIN_CFG, gpu_ids=[local], precision=precision))
    model.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    model.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    weak = scaling == "weak"
    global_batch = args.batch * world if weak else args.batch
    full = synth.synth_train_batch(global_batch, seed=11)
    mine = {k: v.to(dev) for k, v in dataparallel.shard(full, rank, world).items()}
    host = {k: v.pin_memory() for k, v in dataparallel.shard(full, rank, world).items()}
    captured = None
    if graph:
        from dex_gan_grasp_b200.graph_step import CapturedTrainStep
        captured = CapturedTrainStep(model, mine, with_evaluator=True, is_train_gen=True, warmup=2)

    def step(src):
        if captured is not None:  # static input buffers <- this step's batch (device copy, or H2D from pinned memory), one graph launch
            l = captured(src)
            return l, l
        b = {k: v.clone() if src is mine else v.to(dev, non_blocking=True) for k, v in src.items()}
        l1 = model.update_dexgangrasp(b, is_train_gen=True)
        l2 = model.update_dexevaluator(b)
        return l1, l2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step(mine)
    barrier()
    n0 = L.dexgg_launch_count()
    c0 = dataparallel.collective_counters()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in ev:
        e0.record()
        step(mine)
        e1.record()
    barrier()
    launches = (L.dexgg_launch_count() - n0) // max(steps, 1)
    c1 = dataparallel.collective_counters()
    if captured is not None:  # replays do not pass through the host counters: what the graph holds per replay
        launches = captured.kernels_per_replay
        coll = dict(captured.collectives_per_replay)
    else:
        coll = {k: (c1[k] - c0[k]) // max(steps, 1) for k in c1}
    st = stats([a.elapsed_time(b) for a, b in ev])
    t0 = time.perf_counter()
    n_e2e = max(min(steps, 10), 1)
    for _ in range(n_e2e):
        l1, l2 = step(host)
        loss_host = float(l1["total_loss_gen"]) + float(l2["total_loss_eva"])  # device -> host read of the step's result
    torch.cuda.synchronize()
    ms_e2e = 1e3 * (time.perf_counter() - t0) / n_e2e
    t = torch.tensor([st["median"], ms_e2e, st["p90"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    pk, pk_src = peaks()
    tflop = global_batch * (TRAIN_FLOP_PER_ROW_GAN + TRAIN_FLOP_PER_ROW_EVA) / 1e12
    achieved = tflop / (ms * 1e-3) / world  # per GPU
    rec = None
    if rank == 0:
        rec = {"metric": "samples/sec (DexGenerator + Discriminator + DexEvaluator training step)",
               "value": global_batch / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms, "ms_per_step_stats": st, "p90_ms_max_over_ranks": float(t[2]),
               "steps": steps, "warmup": warmup, "scaling": scaling, "global_batch": global_batch, "rows_per_gpu": global_batch // world,
               "precision": precision,
               "dtype": "f32 CUDA-core GEMMs" if precision == "fp32" else "bf16 tensor-core GEMMs / f32 accumulate, f32 master weights",
               "issue": "eager (one Python call per kernel)" if captured is None else "one captured CUDA graph per step (graph_step.CapturedTrainStep)",
               "engine": getattr(model, "train_engine_name", lambda: "autograd Functions over the dexgg kernels (train_ops.py)")(),
               "gpu_launches": int(launches), "collectives_per_step": coll,
               "roofline": {"bound": "tensor", "algorithmic_tflop_per_step": tflop, "achieved": achieved, "unit": "TFLOP/s per GPU",
                            "peak": pk["bf16_tflops_sustained"], "frac": achieved / pk["bf16_tflops_sustained"],
                            "peak_source": pk_src + " bf16_tflops_sustained (step timed inside a long back-to-back run)",
                            "frac_of_burst": achieved / pk["bf16_tflops"]},
               "e2e": {"value": global_batch / (float(t[1]) * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": h2d * world,
                       "d2h_bytes_per_step": 8, "ms_per_step": float(t[1])},
               "last_loss": loss_host}
    if captured is not None:
        captured.close()
    train_ops.set_precision("fp32")
    del model
    torch.cuda.empty_cache()
    return rec


def run_train(args, rank, world, local):
    import torch.distributed as dist
    from dex_gan_grasp_b200 import _native
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _native.check(_native.lib().dexgg_device_check())
    sampler = ClockSampler(local)
    sampler.start()
    rec = train_measure(args, rank, world, local, args.precision, args.steps, max(args.warmup, 3), args.train_scaling, not args.no_graph)
    clocks = sampler.stop()
    if rank == 0:
        line = dict(rec)
        line.update({"n_gpus": world, "higher_is_better": True, "vs_baseline": None, "data": "synthetic", "clocks": clocks,
                     "config": {"workload": "C5 GAN + evaluator training step", "global_batch": rec["global_batch"],
                                "rows_per_gpu": rec["rows_per_gpu"], "issue": rec["issue"], "precision": rec["precision"],
                                "optimizer": "Adam lr 1e-4 wd 5e-4 (fused multi-tensor kernel)",
                                "l2": "working set (activations + weights + Adam state, > 500 MB) exceeds L2"}})
        print(json.dumps(line), flush=True)
    _teardown(world)


def _teardown(world):
    if world > 1:
        import torch.distributed as dist
        # the line is out; a communicator teardown that does not return (seen with captured collectives) must not hold the job
        threading.Thread(target=lambda: (time.sleep(20), os._exit(0)), daemon=True).start()
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- our arm
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    if args.workload == "train":
        run_train(args, rank, world, local)
        return
    import torch.distributed as dist
    from dex_gan_grasp_b200 import _native, dataparallel, synth, tolerances
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    from dex_gan_grasp_b200.pipeline import GraspPipeline

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = _native.lib()
    _native.check(L.dexgg_device_check())
    dev = torch.device("cuda", local)
    B, S = args.clouds, args.samples
    M = B * S

    cfg = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[local], model="dexgangrasp", is_wgan=False,
               precision=args.precision)
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    model = DexGanGrasp(cfg)
    model.DexGANGrasp.load_state_dict(sd_gan)
    model.DexEvaluator.load_state_dict(sd_eva)
    # one captured CUDA graph per step (GraspPipeline use_graph): at 1024 / N clouds per GPU the ~9 launches + allocations that
    # Python would issue per step cost as much host time as the GPU needs for the step
    pipe = GraspPipeline(model.DexGANGrasp, model.DexEvaluator, k=TOPK, thresh=THRESH, z_offset=0.025, use_graph=True)

    # weak workload: rank r owns global objects [r*B, (r+1)*B); strong: rank r owns shard_objects(B) of the SAME B objects
    ids_weak = list(range(rank * B, (rank + 1) * B))
    clouds_np = object_clouds(ids_weak, N_POINTS)
    clouds_h = torch.from_numpy(clouds_np).pin_memory()
    z_h = object_latents(ids_weak, S).pin_memory()
    clouds_d, z_d = clouds_h.to(dev), z_h.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def time_steps(fn, steps, warmup):
        """`steps` timed calls of fn(), each bracketed by its own event pair on the launching stream, an L2 flush (outside the
        pair) before every one; barrier + synchronize on both sides of the timed region."""
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        out = None
        for e0, e1 in ev:
            flush.fill_(1)
            e0.record()
            out = fn()
            e1.record()
        barrier()
        return [a.elapsed_time(b) for a, b in ev], out

    # ---- device-resident throughput (`value`) ----
    # the nvidia-smi poller attaches to the driver when it starts, which stalls the GPU for a moment: start it before the
    # warm-up so that the timed steps only see its 100 ms polling
    sampler = ClockSampler(local)
    sampler.start()
    n_warm = max(args.warmup, 3)
    for _ in range(n_warm):
        pipe.run_device(clouds_d, z_d, S)
    barrier()
    n0 = L.dexgg_launch_count()
    times, out = time_steps(lambda: pipe.run_device(clouds_d, z_d, S), args.steps, 0)
    launches = (L.dexgg_launch_count() - n0) // max(args.steps, 1)
    if launches == 0:  # graph replays do not pass through the library's host-side launch counter: count one eager step
        n1 = L.dexgg_launch_count()
        pipe._run_eager(clouds_d, z_d, S)
        torch.cuda.synchronize()
        launches = L.dexgg_launch_count() - n1
    st_dev = stats(times)
    ms = st_dev["median"]
    kept = int(out["count"].sum().item())
    # the same step repeated for >= 1 s when K steps were less than that (a 0.11 s timed region)
    long_run = None
    if ms * args.steps < 1000.0:
        n_long = int(min(max(1000.0 / ms, args.steps), 400))
        tl, _ = time_steps(lambda: pipe.run_device(clouds_d, z_d, S), n_long, 0)
        long_run = stats(tl)
    clocks = sampler.stop()

    # ---- end to end through the public API with host buffers (`e2e`) ----
    for _ in range(2):
        pipe.run_host(clouds_h, z_h, S)
    barrier()
    t_e2e = []
    for _ in range(min(args.steps, 20)):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res, h2d, d2h = pipe.run_host(clouds_h, z_h, S)
        torch.cuda.synchronize()
        t_e2e.append(time.perf_counter() - t0)
    barrier()
    ms_e2e_sync = 1e3 * float(np.median(t_e2e))  # one step at a time: copy in -> compute -> copy out -> synchronise
    # streaming form of the same public call: K steps, every step's inputs copied from pinned host memory and every step's
    # result read back to the host, the copies of neighbouring steps overlapping the compute (GraspPipeline.run_host_pipelined)
    for _ in pipe.run_host_pipelined(((clouds_h, z_h) for _ in range(2)), S):
        pass
    barrier()
    reps = []
    for _ in range(3):  # three streams of K steps each; the median stream is reported (a stream cannot be split into steps)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_done = 0
        for res, h2d, d2h in pipe.run_host_pipelined(((clouds_h, z_h) for _ in range(args.steps)), S):
            n_done += 1
            kept_e2e = int(res["count"].sum())  # the host-side read of the step's result
        torch.cuda.synchronize()
        reps.append(1e3 * (time.perf_counter() - t0) / args.steps)
    ms_e2e = float(np.median(reps))
    assert n_done == args.steps and kept_e2e == kept, (n_done, kept_e2e, kept)
    barrier()

    # ---- strong split: the SAME 1024 objects sharded by object over the ranks ----
    strong = None
    if not args.no_strong:
        lo, hi = dataparallel.shard_objects(B, rank, world)
        ids_s = list(range(lo, hi))
        cs = torch.from_numpy(object_clouds(ids_s, N_POINTS)).to(dev)
        zs = object_latents(ids_s, S).to(dev)
        ts, out_s = time_steps(lambda: pipe.run_device(cs, zs, S), args.steps, 3)
        st_s = stats(ts)
        kept_s = torch.tensor([float(out_s["count"].sum().item())], device=dev, dtype=torch.float64)
        tt = torch.tensor([st_s["median"], st_s["p90"]], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(kept_s, op=dist.ReduceOp.SUM)
        strong = {"clouds_total": B, "clouds_per_gpu": hi - lo, "ms": float(tt[0]), "p90_ms": float(tt[1]),
                  "grasps_per_s": B * S / (float(tt[0]) * 1e-3), "unit": "grasps/s", "scaling": "strong",
                  "kept_grasps_total": int(kept_s.item()),
                  "note": "objects [lo,hi) per rank via dataparallel.shard_objects, z per object from dataparallel.object_seed "
                          "(kept_grasps_total is the same number at every N)"}
        del cs, zs

    # ---- per-stage device times (isolated, after an L2 flush) and the roofline of the dominant kernels ----
    def timed(fn, n=5):
        ts, r = [], None
        for _ in range(n):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush.fill_(1)
            a.record()
            r = fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        return r, float(np.median(ts))
    enc, t_bps = timed(lambda: pipe.encode(clouds_d))
    (rot, transl, joint, _), t_gen = timed(lambda: model.DexGANGrasp.generator.native_forward(z_d, enc, 0.025, False))
    p, t_eva = timed(lambda: model.DexEvaluator.native_forward(enc, rot, transl))
    k_ = min(TOPK, S)
    f_idx = torch.empty((B, k_), dtype=torch.int32, device=dev)
    f_cnt = torch.empty((B,), dtype=torch.int32, device=dev)
    f_pack = torch.empty((B, k_, 14 + joint.shape[1]), dtype=torch.float32, device=dev)
    _, t_fil = timed(lambda: _native.check(L.dexgg_filter_topk_gather(p.data_ptr(), M, S, THRESH, k_, rot.data_ptr(), transl.data_ptr(),
                                                                      joint.data_ptr(), joint.shape[1], f_idx.data_ptr(), f_cnt.data_ptr(),
                                                                      f_pack.data_ptr(), _native.stream_ptr(dev))))
    stages = {"bps_ms": t_bps, "generator_ms": t_gen, "evaluator_ms": t_eva, "filter_and_gather_ms": t_fil,
              "sum_ms": t_bps + t_gen + t_eva + t_fil,
              "note": "each stage timed alone after an L2 flush (median of 5); the pipelined step keeps rot / transl / scores in L2 "
                      "between stages, so the sum can exceed ms_per_step"}

    pk, pk_src = peaks()
    ms_probe, ops = ctypes.c_float(), ctypes.c_double()
    _native.check(L.dexgg_fp32_pipe_probe(0, 2000, ctypes.byref(ms_probe), ctypes.byref(ops)))
    fp32_tflops = 2 * ops.value / (ms_probe.value * 1e-3) / 1e12  # measured FFMA peak on this GPU, 2 FLOP per lane-op
    mlp_flops = 2.0 * (MAC_PER_GRASP * M + MAC_PER_OBJECT * B)
    mlp_tflops = mlp_flops / ((t_gen + t_eva) * 1e-3) / 1e12
    step_tflops = mlp_flops / (ms * 1e-3) / 1e12
    bps_pairs = 4096.0 * B * N_POINTS
    bps_lane_tops = 3 * bps_pairs / (t_bps * 1e-3) / 1e12  # executed by the filter scan: 3 FFMA per pair (+ exact candidates)
    traffic = None
    tpath = os.path.join(REPO, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        g_, e_ = tj.get("chain_umma_kernel<0>"), tj.get("chain_umma_kernel<2>")
        if g_ and e_:
            traffic = {"file": "profiles/ncu_traffic.json (written by tools/ncu_summary.py traffic from one `ncu --set full` capture of this workload)",
                       "unit": "MB per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
                       "generator": g_["dram_total_mb"], "evaluator": e_["dram_total_mb"], "capture": g_.get("source"),
                       "algorithmic_mb": {"generator": (20 + 96) * M / 1e6, "evaluator": (48 + 4) * M / 1e6,
                                          "note": "z 20 B in + rot/transl/joint 96 B out per grasp; rot/transl 48 B in + score 4 B out; plus the weight stream"}}
    if args.precision == "bf16":
        roof = {"bound": "tensor", "kernel": "chain_umma_kernel<G>,<E> (fused whole-network tcgen05 bf16 kernels; executed-algorithm FLOP of the hoisted chain)",
                "achieved": mlp_tflops, "peak": pk["bf16_tflops"], "unit": "TFLOP/s", "frac": mlp_tflops / pk["bf16_tflops"],
                "peak_source": pk_src + " bf16_tflops (BURST: the two kernels are timed alone, one launch between events after an L2 flush)",
                "frac_of_sustained_peak": mlp_tflops / pk["bf16_tflops_sustained"],
                "whole_step": {"achieved": step_tflops, "frac_of_sustained_peak": step_tflops / pk["bf16_tflops_sustained"],
                               "note": "executed-algorithm MLP FLOP / ms_per_step (BPS, filter and launch gaps included in the time)"},
                "traffic": traffic}
    else:
        roof = {"bound": "fp32", "kernel": "linear_simt_kernel (fp32 CUDA-core mode, all MLP layers)", "achieved": mlp_tflops,
                "peak": fp32_tflops, "unit": "TFLOP/s", "frac": mlp_tflops / fp32_tflops,
                "peak_source": "FFMA probe measured in this run (dexgg_fp32_pipe_probe)", "traffic": None}
    # BPS (second kernel of the step): the filter scan runs on the tensor cores and is read out of TMEM with one 3-input min per two
    # (basis, point) pairs: its bound is the ALU pipe (FMNMX3), measured live by the library's probe (kind 3: dependent FMNMX3 chains)
    ms_alu, ops_alu = ctypes.c_float(), ctypes.c_double()
    _native.check(L.dexgg_fp32_pipe_probe(3, 2000, ctypes.byref(ms_alu), ctypes.byref(ops_alu)))
    alu_tops = ops_alu.value / (ms_alu.value * 1e-3) / 1e12  # T lane-op/s of the ALU pipe on this GPU
    bps_alu_tops = 0.5 * bps_pairs / (t_bps * 1e-3) / 1e12   # algorithmic: one FMNMX3 lane-op per two pairs
    roof["bps_kernel"] = {"bound": "alu", "kernel": "bps_nn_umma_kernel (a() = |x|^2 - 2 b.x by tcgen05 K = 32 split-bf16 MMAs, min scan out of "
                                                    "TMEM, exact direct-form pass on the candidate chunks)",
                          "achieved": bps_alu_tops, "peak": alu_tops,
                          "unit": "T lane-op/s (ALGORITHMIC: one FMNMX3 per two pairs; the scan executes ~0.75 ALU lane-ops per pair "
                                  "with the candidate rule)", "frac": bps_alu_tops / alu_tops,
                          "pairs_per_s": bps_pairs / (t_bps * 1e-3),
                          "fp32_pipe_equivalent": {"lane_tops_if_3_ffma_per_pair": bps_lane_tops, "ffma_peak": fp32_tflops / 2,
                                                   "frac": bps_lane_tops / (fp32_tflops / 2),
                                                   "note": "what the FP32-pipe filter kernel (flag DEXGG_BPS_FMA, r02j: 0.63) would need "
                                                           "for the same pairs/s; > its measured fraction = the gain of the tensor-core scan"},
                          "note": "stage time includes memset + basis pack + finish kernels; profiles/r02n_bps_umma_ncu_full.md holds the pipe "
                                  "utilisation of the kernel alone"}

    # ---- max over ranks ----
    t = torch.tensor([ms, ms_e2e, ms_e2e_sync, st_dev["p90"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, ms_e2e_max, ms_e2e_sync_max = float(t[0]), float(t[1]), float(t[2])

    # ---- parity of the timed workload's first 16 objects against the fp32 oracle pipeline (rank 0, outside all timing) ----
    parity = None
    if rank == 0 and not args.no_parity:
        from dex_gan_grasp_b200 import bps as bps_mod
        from oracle import pipeline_check as pc  # the checker: never timed, never on the product path
        n_chk = min(16, B)
        out_c = pipe.run_device(clouds_d[:n_chk].contiguous(), z_d[:n_chk * S].contiguous(), S)
        torch.cuda.synchronize()
        ref = pc.oracle_pipeline(sd_gan, sd_eva, np.load(bps_mod.BASIS_PATH), clouds_np[:n_chk], z_h[:n_chk * S], S, THRESH, TOPK)
        eps = tolerances.BF16_PIPELINE_EPS if args.precision == "bf16" else 4 * tolerances.FP32_TOL
        parity = pc.check_pipeline(out_c, ref, S, THRESH, TOPK, eps)
        parity["passed"] = bool(parity["bps_bit_exact"] and parity["max_dp"] < eps and parity["set_mismatches"] == 0)
        parity["against"] = "oracle/pipeline_check.py: C BPS -> torch-CPU fp32 generator -> rot6d/offset -> evaluator -> per-object sort/threshold/top-10"

    fp32_line = None
    if args.precision == "bf16" and not args.no_fp32_line:
        # secondary: the 1e-4 parity contract on the TENSOR cores (fp32_tc: bf16 hi + lo pairs, 3 products per MAC, layer by layer)
        # on a quarter of the clouds, same pipeline; and the CUDA-core fp32 mode beside it
        model32 = DexGanGrasp(dict(cfg, precision="fp32_tc"))
        model32.DexGANGrasp.load_state_dict(sd_gan)
        model32.DexEvaluator.load_state_dict(sd_eva)
        pipe32 = GraspPipeline(model32.DexGANGrasp, model32.DexEvaluator, k=TOPK, thresh=THRESH, z_offset=0.025)
        Bq = max(B // 4, 1)
        c32, z32 = clouds_d[:Bq].contiguous(), z_d[:Bq * S].contiguous()
        pipe32.run_device(c32, z32, S)
        torch.cuda.synchronize()
        a32, b32 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a32.record()
        pipe32.run_device(c32, z32, S)
        b32.record()
        b32.synchronize()
        ms32 = a32.elapsed_time(b32)
        tf32 = 3 * 2.0 * (MAC_PER_GRASP * Bq * S + MAC_PER_OBJECT * Bq) / (ms32 * 1e-3) / 1e12  # 3 bf16 products per MAC are EXECUTED
        fp32_line = {"value": Bq * S / (ms32 * 1e-3), "unit": "grasps/s per GPU", "ms": ms32, "sample": f"{Bq} clouds x {S} z",
                     "mode": "fp32_tc: within 1e-4 of the reference on the tensor cores (bf16 hi + lo pairs, hi*hi + lo*hi + hi*lo, fp32 skip branch)",
                     "roofline": {"bound": "tensor", "executed_tflops": tf32, "peak": pk["bf16_tflops_sustained"],
                                  "frac": tf32 / pk["bf16_tflops_sustained"], "frac_of_useful_work": tf32 / 3 / pk["bf16_tflops_sustained"],
                                  "note": "whole pipeline step (BPS and filter included in the time); useful work = executed / 3"}}
        del model32, pipe32
        mcc = DexGanGrasp(dict(cfg, precision="fp32"))
        mcc.DexGANGrasp.load_state_dict(sd_gan)
        mcc.DexEvaluator.load_state_dict(sd_eva)
        pcc = GraspPipeline(mcc.DexGANGrasp, mcc.DexEvaluator, k=TOPK, thresh=THRESH, z_offset=0.025)
        Bc = max(B // 16, 1)
        cc, zc = clouds_d[:Bc].contiguous(), z_d[:Bc * S].contiguous()
        pcc.run_device(cc, zc, S)
        torch.cuda.synchronize()
        a32.record()
        pcc.run_device(cc, zc, S)
        b32.record()
        b32.synchronize()
        fp32_line["cuda_core_fp32_mode"] = {"value": Bc * S / (a32.elapsed_time(b32) * 1e-3), "unit": "grasps/s per GPU", "sample": f"{Bc} clouds x {S} z"}
        del mcc, pcc

    # ---- C1: one real object, latency through the class surface and through the pipeline API (rank 0) ----
    c1 = None
    if rank == 0 and not args.no_c1:
        try:
            c1 = c1_latency(model, dev)
            if not args.no_cpu_baseline:
                c1["cpu_port"] = c1_cpu_port(400)
        except Exception as e:
            c1 = {"error": repr(e)[:300]}

    # ---- C5 training block (every rank takes part: NCCL) ----
    del pipe, model, clouds_d, z_d, flush
    torch.cuda.empty_cache()
    train = None
    if not args.no_train:
        try:
            train = train_measure(args, rank, world, local, "bf16", args.train_steps, 5, "strong", True)
        except Exception as e:  # the headline line must still be printed
            train = {"error": repr(e)[:300]} if rank == 0 else None

    gpu_base = None
    if rank == 0 and not args.no_gpu_baseline:
        try:
            g, dt = gpu_reference_sample(16, S, N_POINTS, dev)
            gpu_base = {"value": g / dt, "unit": "grasps/s", "kind": "torch-eager fp32 on this GPU (TF32 off): the reference algorithm with stock ATen kernels, "
                        "per-object loop with its np.tile / H2D / D2H round trips (dexgangrasp.py:393-442,495-514)",
                        "sample": f"first 16 objects x {S} z, best of 3, {dt:.3f} s", "seconds": dt}
        except Exception as e:
            gpu_base = {"error": repr(e)[:300]}
    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        n_cpu = 128
        g, dt, cores = cpu_reference_sample(n_cpu, S, N_POINTS, 0)
        cpu = {"value": g / dt, "unit": "grasps/s", "cores": cores, "kind": "port",
               "sample": f"first {n_cpu} objects x {S} z of the same workload, oracle/ (C BPS + torch-CPU networks), {dt:.2f} s"}
    if rank == 0:
        line = {
            "metric": METRIC, "value": world * M / (ms_max * 1e-3), "unit": "grasps/s", "n_gpus": world, "steps": args.steps,
            "warmup": n_warm, "ms_per_step": ms_max, "ms_per_step_stats": st_dev, "p90_ms_max_over_ranks": float(t[3]),
            "long_run": long_run, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (BPS) + " + ("bf16 tensor-core MMA / f32 accumulate (MLPs)" if args.precision == "bf16" else "f32 CUDA-core (MLPs)"),
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "clouds_per_gpu": B, "points": N_POINTS, "samples_per_cloud": S, "topk": TOPK,
                       "thresh": THRESH, "precision": args.precision, "l2": "256 MB flush write between timed iterations",
                       "timing": "per-step CUDA events, median over K steps, max over ranks",
                       "issue": "one captured CUDA graph per step (GraspPipeline use_graph=True)",
                       "weights": "seeded synthetic state dicts (no trained checkpoints exist offline)"},
            "e2e": {"value": world * M / (ms_e2e_max * 1e-3), "unit": "grasps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e_max, "streams_ms_per_step": [round(r, 4) for r in reps], "mode": "streaming (median of 3 streams of K steps): K steps through GraspPipeline.run_host_pipelined (H2D of step i+1 and D2H of "
                    "step i-1 overlap the compute of step i; every step's inputs come from pinned host memory, every result is read on the host)",
                    "one_step_at_a_time": {"value": world * M / (ms_e2e_sync_max * 1e-3), "ms_per_step": ms_e2e_sync_max}},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "stages": stages, "parity_check": parity,
            "strong": strong, "train": train, "c1_latency": c1, "cpu_baseline": cpu, "gpu_baseline": gpu_base,
            "kept_grasps_last_step": kept, "fp32_exact_mode": fp32_line,
        }
        print(json.dumps(line), flush=True)
    _teardown(world)


if __name__ == "__main__":
    main()
