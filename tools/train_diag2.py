import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dex_gan_grasp_b200 import synth
from oracle import nets_oracle as no
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
SEED, B = 7, 192
CFG = dict(latentD=5, n_hand_joints=12, is_train=True, gpu_ids=[0], model="dexgangrasp", is_wgan=False, bce_weight=1.0,
           train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
           lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
           init_gain=0.02)
batch = synth.synth_train_batch(B, seed=11)
g = torch.Generator().manual_seed(5)
masks = [(torch.rand(B, 512, generator=g) >= 0.3).float() / 0.7 for _ in range(3)]
sd32 = {k: v.clone() for k, v in synth.synth_eva_state_dict(SEED).items()}
_, g32 = no.evaluator_step(sd32, {}, copy.deepcopy(batch), dropout_masks=masks)
junk = []
for it in range(6):
    m = DexGanGrasp(dict(CFG))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED)); m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    m.DexEvaluator.dropout_masks = [x.cuda() for x in masks]
    m.update_dexevaluator(copy.deepcopy(batch))
    torch.cuda.synchronize()
    named = dict(m.DexEvaluator.named_parameters())
    worst = sorted(((float((named[k].grad.cpu() - g32[k]).abs().max()) / max(1e-12, float(g32[k].abs().max())), k) for k in g32 if not k.endswith(("fc1.bias", "fc2.bias"))), reverse=True)[:4]
    print(it, " ".join("%s=%.2e" % (k, e) for e, k in worst), flush=True)
    # churn the allocator so later iterations reuse dirty memory
    junk = [torch.randn(1 << 20, device="cuda") * 1e3 for _ in range(8)]
    del junk
