"""One C3-sized BPS encode (1024 clouds x 2048 points) - the target of `ncu -k regex:bps_nn` captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dex_gan_grasp_b200 import bps as bps_mod
flags = int(os.environ.get("BPS_FLAGS", "0"))
basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
pts = 0.08 * torch.randn((1024, 2048, 3), device="cuda")
for _ in range(3):
    d, i = bps_mod.bps_encode_device(basis, pts, flags=flags)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); d, i = bps_mod.bps_encode_device(basis, pts, flags=flags); e1.record(); torch.cuda.synchronize()
print("bps ms", e0.elapsed_time(e1))
