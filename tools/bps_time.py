"""BPS stage alone at the C3 shape (1024 clouds x 2048 points x 4096 basis points): default tensor-core scan against the
FP32-pipe filter kernel (flag DEXGG_BPS_FMA) and the direct kernel; CUDA events, L2 flushed between launches."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dex_gan_grasp_b200 import bps as bps_mod

B, N = int(os.environ.get("BPS_B", 1024)), int(os.environ.get("BPS_N", 2048))
basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
g = torch.Generator(device="cuda").manual_seed(0)
pts = 0.08 * torch.randn((B, N, 3), device="cuda", generator=g)
if os.environ.get("BPS_SYNTH", "0") == "1":  # the benchmark's single-view surface clouds
    from dex_gan_grasp_b200 import synth
    pts = torch.from_numpy(synth.synth_clouds(B, N)).cuda()
flush = torch.empty((256 << 20,), dtype=torch.uint8, device="cuda")
out = {}
ref = None
for name, flags in (("umma", 0), ("fma", 4), ("direct", 2)):
    ts = []
    for it in range(8):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        d, i = bps_mod.bps_encode_device(basis, pts, flags=flags)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    if ref is None:
        ref = (d, i)
    out[name] = {"ms_median": float(np.median(ts[2:])), "ms_min": float(min(ts[2:])), "equal_to_first": bool(torch.equal(d, ref[0]) and torch.equal(i, ref[1]))}
print(json.dumps({"synth_surfaces": os.environ.get("BPS_SYNTH", "0") == "1", "clouds": B, "points": N, "basis": int(basis.shape[0]), "stage_ms": out}))
