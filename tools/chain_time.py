"""C3-sized generator + evaluator forwards (1024 objects x 1000 z): CUDA-event time of each fused chain kernel call, L2 flushed."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dex_gan_grasp_b200 import synth
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
cfg = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False, precision="bf16")
m = DexGanGrasp(cfg)
m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(7)); m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(7))
B, S = 1024, 1000
bps = (0.1 + 0.03 * torch.randn(B, 4096)).cuda()
z = torch.randn(B * S, 5).cuda()
flush = torch.empty((256 << 20,), dtype=torch.uint8, device="cuda")
tg, te = [], []
for it in range(8):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    flush.zero_()
    ev[0].record()
    rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    ev[1].record()
    flush.zero_()
    ev[2].record()
    p = m.DexEvaluator.native_forward(bps, rot, transl)
    ev[3].record()
    torch.cuda.synchronize()
    tg.append(ev[0].elapsed_time(ev[1])); te.append(ev[2].elapsed_time(ev[3]))
print(json.dumps({"generator_ms": float(np.median(tg[2:])), "evaluator_ms": float(np.median(te[2:])),
                  "checksum": [float(p.double().sum()), float(transl.double().abs().sum())]}))
