#!/usr/bin/env python
"""bench.py — grasps/sec of the DexGANGrasp hot path (BPS -> DexGenerator -> DexEvaluator -> top-k) on B200.

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): 1024 synthetic single-view point clouds of
2048 points x 1000 latent samples per cloud -> evaluator -> per-object top-10 above threshold 0.5.  One process per GPU;
every rank runs the full workload on its own objects (no data-path collective) -> "scaling": "weak".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--precision fp32|bf16] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...

`--impl reference` times the CPU restatement of the reference path (oracle/: C BPS + torch-CPU networks, all host cores)
on a bounded sample of the same workload.  Prints ONE JSON line on rank 0.
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

N_CLOUDS, N_POINTS, N_SAMPLES, TOPK, THRESH, SEED = 1024, 2048, 1000, 10, 0.5, 7
MAC_PER_GRASP, MAC_PER_OBJECT = 1_881_600, 15_728_640  # hoisted algorithm (SURVEY.md §8a note H)
METRIC = "grasps/sec (BPS->DexGenerator->DexEvaluator->top-k)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("DEXGG_PRECISION", "bf16"), choices=["fp32", "bf16"],
                    help="MLP arithmetic: bf16 = tcgen05 tensor cores (headline, north_star), fp32 = CUDA-core 1e-4 mode")
    ap.add_argument("--no-fp32-line", action="store_true", help="skip the secondary fp32-exact-mode measurement")
    ap.add_argument("--clouds", type=int, default=N_CLOUDS)
    ap.add_argument("--samples", type=int, default=N_SAMPLES)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c3", choices=["c3", "train"],
                    help="c3 = the headline inference metric; train = secondary C5 line (GAN + evaluator training step, global batch 4096)")
    ap.add_argument("--batch", type=int, default=4096, help="global training batch (C5)")
    ap.add_argument("--train-scaling", default="strong", choices=["strong", "weak"],
                    help="C5: strong = the global batch is split over the GPUs (SURVEY 8e); weak = --batch rows on every GPU")
    ap.add_argument("--no-graph", action="store_true", help="C5: issue the step eagerly instead of replaying the captured CUDA graph")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# --------------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference_sample(n_objects, n_samples, n_points, seed0, repeats=1):
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
    clouds = [synth.synth_cloud(n_points, seed0 + i) for i in range(n_objects)]
    zs = [synth.synth_latents(n_samples, 5, seed=seed0 + i) for i in range(n_objects)]
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        kept = 0
        for c, z in zip(clouds, zs):
            d = np.empty(4096, np.float32)
            idx = np.empty(4096, np.int32)
            lib.dexgg_oracle_bps_encode_mt(vp(basis), 4096, vp(c), len(c), vp(d), vp(idx), min(cores, 256))
            bps_t = torch.from_numpy(np.tile(d, (n_samples, 1)))
            g = no.generate_poses(sd_gan, z, bps_t, 0.025)
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
    n_obj = 8
    times = []
    for i in range(args.warmup + args.steps):
        g, dt, cores = cpu_reference_sample(n_obj, args.samples, N_POINTS, 1000 + 17 * i)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    val = n_obj * args.samples / (ms * 1e-3)
    sample = f"{n_obj} objects x {args.samples} z per step (of {args.clouds} x {args.samples})"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "grasps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C3 batched inference: synthetic clouds (2048 pts) x 1000 latent samples -> evaluator top-10",
                       "clouds": args.clouds, "points": N_POINTS, "samples_per_cloud": args.samples, "topk": TOPK, "thresh": THRESH},
            "cpu_baseline": {"value": val, "unit": "grasps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "grasps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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


# --------------------------------------------------------------------------------------------- C5: training step (secondary)
TRAIN_CFG = dict(latentD=5, n_hand_joints=12, is_train=True, model="dexgangrasp", is_wgan=False, bce_weight=1.0,
                 train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
                 lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
                 init_gain=0.02)


def run_train(args, rank, world, local):
    """C5 (BASELINE.json configs[4]): one step = DexGanGrasp.update_dexgangrasp (D step + G step) + update_dexevaluator on a
    global batch of 4096 synthetic rows, sharded 4096/N per rank, gradients and BatchNorm statistics allreduced over NCCL
    (train.py:233-260 drives the same two calls per batch).  Strong scaling: the global batch is fixed."""
    import torch.distributed as dist
    from dex_gan_grasp_b200 import _native, dataparallel, synth
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = _native.lib()
    _native.check(L.dexgg_device_check())
    dev = torch.device("cuda", local)
    model = DexGanGrasp(dict(TRAIN_CFG, gpu_ids=[local], precision=args.precision))
    model.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    model.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    weak = args.train_scaling == "weak"
    global_batch = args.batch * world if weak else args.batch
    full = synth.synth_train_batch(global_batch, seed=11)
    mine = {k: v.to(dev) for k, v in dataparallel.shard(full, rank, world).items()}
    host = {k: v.pin_memory() for k, v in dataparallel.shard(full, rank, world).items()}
    captured = None
    if not args.no_graph:
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

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        step(mine)
    barrier()
    n0 = L.dexgg_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step(mine)
    e1.record()
    barrier()
    launches = (L.dexgg_launch_count() - n0) // max(args.steps, 1)
    if captured is not None:  # replays do not pass through the host launch counter: the graph's own kernel-node count
        launches = captured.kernels_per_replay
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1) / args.steps
    t0 = time.perf_counter()
    for _ in range(args.steps):
        l1, l2 = step(host)
        loss_host = float(l1["total_loss_gen"]) + float(l2["total_loss_eva"])  # device -> host read of the step's result
    torch.cuda.synchronize()
    ms_e2e = 1e3 * (time.perf_counter() - t0) / args.steps
    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    if rank == 0:
        line = {"metric": "samples/sec (DexGenerator + Discriminator + DexEvaluator training step)", "value": global_batch / (float(t[0]) * 1e-3),
                "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(t[0]),
                "higher_is_better": True, "scaling": args.train_scaling, "vs_baseline": None,
                "dtype": "f32 CUDA-core GEMMs" if args.precision == "fp32" else "bf16 tensor-core GEMMs / f32 accumulate, f32 master weights",
                "data": "synthetic",
                "config": {"workload": "C5 GAN + evaluator training step", "global_batch": global_batch, "rows_per_gpu": global_batch // world,
                           "issue": "eager (one Python call per kernel)" if captured is None else "one captured CUDA graph per step (graph_step.CapturedTrainStep)",
                           "precision": args.precision, "optimizer": "Adam lr 1e-4 wd 5e-4 (fused multi-tensor kernel)",
                           "l2": "working set (activations + weights + Adam state, > 500 MB) exceeds L2"},
                "e2e": {"value": global_batch / (float(t[1]) * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": h2d * world,
                        "d2h_bytes_per_step": 8, "ms_per_step": float(t[1])},
                "gpu_launches": int(launches), "clocks": clocks, "last_loss": loss_host}
        print(json.dumps(line), flush=True)
    if captured is not None:
        captured.close()
    if world > 1:
        # the line is out; a communicator teardown that does not return (seen with captured collectives) must not hold the job
        import threading
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
    if args.workload == "train":
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
        run_train(args, rank, world, local)
        return
    import torch.distributed as dist
    from dex_gan_grasp_b200 import _native, synth
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    from dex_gan_grasp_b200.pipeline import GraspPipeline

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
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
    model = DexGanGrasp(cfg)
    model.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    model.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    pipe = GraspPipeline(model.DexGANGrasp, model.DexEvaluator, k=TOPK, thresh=THRESH, z_offset=0.025)

    clouds_h = torch.from_numpy(synth.synth_clouds(B, N_POINTS, seed0=1000 + 100000 * rank)).pin_memory()
    z_h = torch.randn((M, 5), generator=torch.Generator().manual_seed(1234 + rank)).pin_memory()
    clouds_d, z_d = clouds_h.to(dev), z_h.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (`value`) ----
    # the nvidia-smi poller attaches to the driver when it starts, which stalls the GPU for a moment: start it before the
    # warm-up so that the timed steps only see its 100 ms polling
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        pipe.run_device(clouds_d, z_d, S)
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    n0 = L.dexgg_launch_count()
    for e0, e1 in ev:
        flush.fill_(1)  # L2 flush between timed iterations (outside the event pair)
        e0.record()
        out = pipe.run_device(clouds_d, z_d, S)
        e1.record()
    barrier()
    launches = (L.dexgg_launch_count() - n0) // max(args.steps, 1)
    clocks = sampler.stop()
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    kept = int(out["count"].sum().item())

    # ---- end to end through the public API with host buffers (`e2e`) ----
    for _ in range(2):
        pipe.run_host(clouds_h, z_h, S)
    barrier()
    t_e2e = []
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res, h2d, d2h = pipe.run_host(clouds_h, z_h, S)
        torch.cuda.synchronize()
        t_e2e.append(time.perf_counter() - t0)
    barrier()
    ms_e2e_sync = 1e3 * float(np.mean(t_e2e))  # one step at a time: copy in -> compute -> copy out -> synchronise
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

    # ---- per-stage device times (one instrumented step) and the roofline of the dominant kernel ----
    def timed(fn):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.fill_(1)
        a.record()
        r = fn()
        b.record()
        b.synchronize()
        return r, a.elapsed_time(b)
    enc, t_bps = timed(lambda: pipe.encode(clouds_d))
    (rot, transl, joint, _), t_gen = timed(lambda: model.DexGANGrasp.generator.native_forward(z_d, enc, 0.025, False))
    p, t_eva = timed(lambda: model.DexEvaluator.native_forward(enc, rot, transl))
    stages = {"bps_ms": t_bps, "generator_ms": t_gen, "evaluator_ms": t_eva, "filter_and_gather_ms": max(ms - t_bps - t_gen - t_eva, 0.0)}

    pk, pk_src = peaks()
    ms_probe, ops = ctypes.c_float(), ctypes.c_double()
    _native.check(L.dexgg_fp32_pipe_probe(0, 2000, ctypes.byref(ms_probe), ctypes.byref(ops)))
    fp32_tflops = 2 * ops.value / (ms_probe.value * 1e-3) / 1e12  # measured FFMA peak on this GPU, 2 FLOP per lane-op
    mlp_flops = 2.0 * (MAC_PER_GRASP * M + MAC_PER_OBJECT * B)
    mlp_tflops = mlp_flops / ((t_gen + t_eva) * 1e-3) / 1e12
    bps_pairs = 4096.0 * B * N_POINTS
    bps_lane_tops = 3 * bps_pairs / (t_bps * 1e-3) / 1e12  # executed by the filter scan: 3 FFMA per pair (+ exact candidates)
    if args.precision == "bf16":
        roof = {"bound": "tensor", "kernel": "chain_umma_kernel<G>,<E> (fused whole-network tcgen05 bf16 kernels; executed-algorithm FLOP of the hoisted chain)", "achieved": mlp_tflops,
                "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s", "frac": mlp_tflops / pk["bf16_tflops_sustained"],
                "peak_source": pk_src + " bf16_tflops_sustained",
                "traffic": {"unit": "MB per launch (ncu dram__bytes_read+write, profiles/r01l_chain_ncu_full.md)", "generator": 69.8, "evaluator": 62.4,
                            "algorithmic": "generator 20 B in (z) + 96 B out per grasp = 118.8 MB + weights; evaluator 48 B in + 4 B out = 53.2 MB + weights: "
                                           "measured DRAM traffic is BELOW the algorithmic bytes: part of z / rot / transl / joint is still resident in the 126 MB L2 when the kernel ends"}}
    else:
        roof = {"bound": "fp32", "kernel": "linear_simt_kernel (fp32 CUDA-core mode, all MLP layers)", "achieved": mlp_tflops,
                "peak": fp32_tflops, "unit": "TFLOP/s", "frac": mlp_tflops / fp32_tflops,
                "peak_source": "FFMA probe measured in this run (dexgg_fp32_pipe_probe)", "traffic": None}
    roof["bps_kernel"] = {"bound": "fp32", "kernel": "bps_nn_filter_kernel (expanded-form filter scan + exact direct-form candidates)",
                          "achieved": bps_lane_tops, "peak": fp32_tflops / 2,
                          "unit": "T lane-op/s (3 FFMA per pair executed by the scan)", "frac": bps_lane_tops / (fp32_tflops / 2),
                          "direct_form_equivalent_frac": 2 * bps_lane_tops / (fp32_tflops / 2),
                          "note": "SURVEY 8d counts 6 FP32-pipe ops per pair for the direct-difference scan; the filter needs 3, so the "
                                  "same pairs/s is reported both ways (executed ops, and what a direct-form scan would need)"}

    # ---- max over ranks ----
    t = torch.tensor([ms, ms_e2e, ms_e2e_sync], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, ms_e2e_max, ms_e2e_sync_max = float(t[0]), float(t[1]), float(t[2])

    fp32_line = None
    if args.precision == "bf16" and not args.no_fp32_line:
        # secondary: the fp32 CUDA-core mode (1e-4 parity contract) on a quarter of the clouds, same pipeline
        model32 = DexGanGrasp(dict(cfg, precision="fp32"))
        model32.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
        model32.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
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
        fp32_line = {"value": Bq * S / (a32.elapsed_time(b32) * 1e-3), "unit": "grasps/s per GPU", "ms": a32.elapsed_time(b32),
                     "sample": f"{Bq} clouds x {S} z", "note": "fp32 CUDA-core MLP mode (within 1e-4 of the reference)"}
        del model32, pipe32

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        g, dt, cores = cpu_reference_sample(16, S, N_POINTS, 1000)
        cpu = {"value": g / dt, "unit": "grasps/s", "cores": cores, "kind": "port",
               "sample": f"first 16 objects x {S} z of the same workload, oracle/ (C BPS + torch-CPU networks), {dt:.2f} s"}
    if rank == 0:
        line = {
            "metric": METRIC, "value": world * M / (ms_max * 1e-3), "unit": "grasps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (BPS) + " + ("bf16 tensor-core MMA / f32 accumulate (MLPs)" if args.precision == "bf16" else "f32 CUDA-core (MLPs)"),
            "data": "synthetic",
            "config": {"workload": "C3 batched inference: synthetic clouds (2048 pts) x 1000 latent samples -> evaluator top-10",
                       "clouds_per_gpu": B, "points": N_POINTS, "samples_per_cloud": S, "topk": TOPK, "thresh": THRESH,
                       "precision": args.precision, "l2": "256 MB flush write between timed iterations",
                       "weights": "seeded synthetic state dicts (no trained checkpoints exist offline)"},
            "e2e": {"value": world * M / (ms_e2e_max * 1e-3), "unit": "grasps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e_max, "streams_ms_per_step": [round(r, 4) for r in reps], "mode": "streaming (median of 3 streams of K steps): K steps through GraspPipeline.run_host_pipelined (H2D of step i+1 and D2H of "
                    "step i-1 overlap the compute of step i; every step's inputs come from pinned host memory, every result is read on the host)",
                    "one_step_at_a_time": {"value": world * M / (ms_e2e_sync_max * 1e-3), "ms_per_step": ms_e2e_sync_max}},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "stages": stages, "cpu_baseline": cpu,
            "kept_grasps_last_step": kept, "fp32_exact_mode": fp32_line,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
