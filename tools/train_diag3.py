import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from dex_gan_grasp_b200 import synth, train_ops as T
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
SEED, B = 7, 192
CFG = dict(latentD=5, n_hand_joints=12, is_train=True, gpu_ids=[0], model="dexgangrasp", is_wgan=False, bce_weight=1.0,
           train_ffhgenerator=True, train_ffhevaluator=True, beta1=0.9, weight_decay=5e-4, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4,
           lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=5, weight_init_type="xavier",
           init_gain=0.02)
torch.backends.cuda.matmul.allow_tf32 = False
batch = {k: v.cuda().float() for k, v in synth.synth_train_batch(B, seed=11).items()}
g = torch.Generator().manual_seed(5)
masks = [((torch.rand(B, 512, generator=g) >= 0.3).float() / 0.7).cuda() for _ in range(3)]
lrelu = lambda t: F.leaky_relu(t, 0.2)
_orig_lb = T.lrelu_backward
def _dbg_lb(g, y):
    out = _orig_lb(g, y)
    torch.cuda.synchronize()
    ref = torch.where(y > 0, g, 0.2 * g)
    e = float((out - ref).abs().max())
    if e > 1e-7 * max(1e-20, float(ref.abs().max())):
        print("  lrelu_backward MISMATCH", e, tuple(g.shape), g.stride(), y.stride(), g.data_ptr() % 256, y.data_ptr() % 256, g.is_contiguous(), flush=True)
    return out
T.lrelu_backward = _dbg_lb
_orig_rl = T.raw_linear
def _dbg_rl(A, ta, W, tw, M, N, K, bias=None, act=0, accumulate=0, out=None):
    o = _orig_rl(A, ta, W, tw, M, N, K, bias, act, accumulate, out)
    torch.cuda.synchronize()
    Am = (A.t() if ta else A)[:M, :K].double(); Wm = (W.t() if tw else W)[:N, :K].double()
    ref = Am @ Wm.t()
    if bias is not None: ref = ref + bias.double()
    if act: ref = F.leaky_relu(ref, 0.2)
    e = float((o.double() - ref).abs().max())
    if e > 1e-4 * max(1e-20, float(ref.abs().max())):
        print("  raw_linear MISMATCH", e, float(ref.abs().max()), (ta, tw, M, N, K), A.stride(), W.stride(), A.data_ptr() % 16, W.data_ptr() % 16, flush=True)
    return o
T.raw_linear = _dbg_rl

def ref_rb(rb, x, keep, name):
    xin = lrelu(F.linear(x, rb.fc3.weight, rb.fc3.bias)); keep[name + ".xin"] = xin
    h1 = F.linear(x, rb.fc1.weight, rb.fc1.bias); keep[name + ".h1"] = h1
    h = lrelu(F.batch_norm(h1, None, None, rb.bn1.weight, rb.bn1.bias, True, 0.1, 1e-5)); keep[name + ".h"] = h
    h2 = F.linear(h, rb.fc2.weight, rb.fc2.bias); keep[name + ".h2"] = h2
    y = F.batch_norm(h2, None, None, rb.bn2.weight, rb.bn2.bias, True, 0.1, 1e-5)
    out = lrelu(xin + y); keep[name + ".out"] = out
    return out

def our_rb(rb, x, keep, name):
    xin = T.linear(x, rb.fc3.weight, rb.fc3.bias, act=True); keep[name + ".xin"] = xin
    h1 = T.linear(x, rb.fc1.weight, rb.fc1.bias); keep[name + ".h1"] = h1
    h = T.batch_norm(h1, rb.bn1, act=True); keep[name + ".h"] = h
    h2 = T.linear(h, rb.fc2.weight, rb.fc2.bias); keep[name + ".h2"] = h2
    out = T.batch_norm(h2, rb.bn2, act=False, residual=xin, final_act=True); keep[name + ".out"] = out
    return out

def run(eva, ours):
    keep = {}
    rbf = our_rb if ours else ref_rb
    X = torch.cat([batch["bps_object"], batch["rot_matrix"].reshape(B, 9), batch["transl"]], dim=1)
    X0 = T.batch_norm(X, eva.bn1) if ours else F.batch_norm(X, None, None, eva.bn1.weight, eva.bn1.bias, True, 0.1, 1e-5)
    keep["X0"] = X0
    X1 = rbf(eva.rb1, X0, keep, "rb1") * masks[0]; keep["X1"] = X1
    X2 = rbf(eva.rb2, torch.cat([X1, X0], 1), keep, "rb2") * masks[1]; keep["X2"] = X2
    X3 = rbf(eva.rb3, torch.cat([X2, X0], 1), keep, "rb3") * masks[2]; keep["X3"] = X3
    lin = T.linear if ours else F.linear
    p = torch.sigmoid(lin(X3, eva.out_success.weight, eva.out_success.bias))
    for v in keep.values(): v.retain_grad()
    loss = F.binary_cross_entropy(p, batch["label"].unsqueeze(-1))
    eva.zero_grad()
    loss.backward()
    return {k: (v.detach().clone(), v.grad.detach().clone()) for k, v in keep.items()}

for it in range(6):
    m = DexGanGrasp(dict(CFG))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    m.DexEvaluator.train()
    r = run(m.DexEvaluator, False)
    o = run(m.DexEvaluator, True)
    torch.cuda.synchronize()
    bad = []
    for k in r:
        ev = float((r[k][0] - o[k][0]).abs().max()) / max(1e-20, float(r[k][0].abs().max()))
        eg = float((r[k][1] - o[k][1]).abs().max()) / max(1e-20, float(r[k][1].abs().max()))
        if ev > 1e-4 or eg > 1e-4: bad.append("%s v=%.1e g=%.1e" % (k, ev, eg))
    print(it, bad if bad else "ok", flush=True)
    junk = [torch.randn(1 << 20, device="cuda") * 1e3 for _ in range(8)]
    del junk
