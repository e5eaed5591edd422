"""Diagnostic: per-tensor gradient error of the CUDA training step vs the CPU oracle, calibrated by the oracle's own
fp32-vs-fp64 difference (rounding-noise floor)."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no
SEED, B = 7, 192
CFG = dict(latentD=5, n_hand_joints=12, is_train=True, gpu_ids=[0], model="dexgangrasp", is_wgan=False, bce_weight=1.0,
           train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
           lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
           init_gain=0.02)
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp

def to64(sd): return {k: (v.double() if v.is_floating_point() else v.clone()) for k, v in sd.items()}
m = DexGanGrasp(dict(CFG))
m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED)); m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
batch = synth.synth_train_batch(B, seed=11)
g = torch.Generator().manual_seed(5)
masks = [(torch.rand(B, 512, generator=g) >= 0.3).float() / 0.7 for _ in range(3)]
sd32 = {k: v.clone() for k, v in synth.synth_eva_state_dict(SEED).items()}
_, g32 = no.evaluator_step(sd32, {}, copy.deepcopy(batch), dropout_masks=masks)
sd64 = to64(synth.synth_eva_state_dict(SEED))
torch.set_default_dtype(torch.float64)
class B64(dict): pass
b64 = {k: v.double() for k, v in batch.items()}
import torch.nn.functional as F
keys = no._params(sd64)
for k in keys: sd64[k].requires_grad_(True)
p = no.evaluator_forward(sd64, b64["bps_object"], b64["rot_matrix"], b64["transl"], training=True, dropout_masks=[x.double() for x in masks])
loss = F.binary_cross_entropy(p, b64["label"].unsqueeze(-1))
g64 = dict(zip(keys, torch.autograd.grad(loss, [sd64[k] for k in keys])))
torch.set_default_dtype(torch.float32)
m.DexEvaluator.dropout_masks = [x.cuda() for x in masks]
m.update_dexevaluator(copy.deepcopy(batch))
named = dict(m.DexEvaluator.named_parameters())
print("%-28s %10s %12s %12s" % ("tensor", "max|g|", "ours-f64", "cpu32-f64"))
for k in keys:
    gm = float(g64[k].abs().max())
    e1 = float((named[k].grad.cpu().double() - g64[k]).abs().max()) / gm
    e2 = float((g32[k].double() - g64[k]).abs().max()) / gm
    print("%-28s %10.3e %12.3e %12.3e" % (k, gm, e1, e2))
