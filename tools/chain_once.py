"""C3-sized generator + evaluator forwards (1024 objects x 1000 z) - the target of `ncu -k regex:chain_umma` captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dex_gan_grasp_b200 import synth
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
cfg = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False, precision="bf16")
m = DexGanGrasp(cfg)
m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(7)); m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(7))
B, S = 1024, 1000
bps = (0.1 + 0.03 * torch.randn(B, 4096)).cuda()
z = torch.randn(B * S, 5).cuda()
for _ in range(3):
    rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    p = m.DexEvaluator.native_forward(bps, rot, transl)
torch.cuda.synchronize()
