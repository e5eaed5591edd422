"""GPU probe: FP32-pipe peaks (FFMA / FFMA2 / FADD2 / FMNMX3) and BPS kernel timings. Writes gpurun_out/probe_bps.json."""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dex_gan_grasp_b200 import _native, bps as bps_mod  # noqa: E402

L = _native.lib()
out = {"gpu": torch.cuda.get_device_name(0), "sms": torch.cuda.get_device_properties(0).multi_processor_count}
names = {0: "ffma", 1: "ffma2", 2: "fadd2", 3: "fmnmx3"}
for kind, nm in names.items():
    ms = ctypes.c_float()
    ops = ctypes.c_double()
    _native.check(L.dexgg_fp32_pipe_probe(kind, 4000, ctypes.byref(ms), ctypes.byref(ops)))
    out["probe_" + nm] = {"ms": ms.value, "lane_ops_per_s": ops.value / (ms.value * 1e-3)}
    print(nm, ms.value, "ms", ops.value / (ms.value * 1e-3) / 1e12, "T lane-ops/s", flush=True)

basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(iters):
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


ws = bps_mod._Workspace()
res = []
# C4 (BASELINE.json configs[3]): N = 1k ... 100k points, B = 1 ... 4096 clouds, B*N <= 2^28
SWEEP = [(1024, 2048), (1, 1024), (1, 16384), (1, 100000), (8, 100000), (64, 16384), (64, 65536), (512, 8192), (512, 32768),
         (4096, 1024), (4096, 4096), (2048, 100000)]
for (B, N) in SWEEP:
    pts = 0.08 * torch.randn((B, N, 3), device="cuda")
    for name, flags in [("umma", 0), ("filter", 4), ("direct", 2), ("scalar", 1)]:
        if name in ("direct", "scalar") and os.environ.get("PROBE_FAST", "1") == "1":
            continue
        if name == "scalar" and B * N > 3e6:
            continue
        med, mn = timeit(lambda: bps_mod.bps_encode_device(basis, pts, flags=flags, ws=ws))
        pairs = 4096.0 * B * N
        r = {"B": B, "N": N, "variant": name, "ms_median": med, "ms_min": mn, "gpairs_per_s": pairs / (med * 1e-3) / 1e9,
             "fp32_lane_ops_per_s_executed": (0 if name == "umma" else 3 if name == "filter" else 6) * pairs / (med * 1e-3),
             "fp32_lane_ops_per_s_direct_form": 6 * pairs / (med * 1e-3)}
        res.append(r)
        print(r, flush=True)
out["bps"] = res
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/probe_bps.json", "w"), indent=1)
