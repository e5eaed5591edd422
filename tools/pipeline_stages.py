"""Back-to-back stage times inside GraspPipeline.run_device (events between the stages, no L2 flush inside the step)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dex_gan_grasp_b200 import synth, _native
from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
from dex_gan_grasp_b200.pipeline import GraspPipeline
cfg = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False, precision="bf16")
m = DexGanGrasp(cfg)
m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(7)); m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(7))
pipe = GraspPipeline(m.DexGANGrasp, m.DexEvaluator, k=10, thresh=0.5, z_offset=0.025)
B, S = 1024, 1000
clouds = torch.from_numpy(synth.synth_clouds(B, 2048, seed0=1000)).cuda()
z = torch.randn(B * S, 5).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
L = _native.lib()
for _ in range(3):
    pipe.run_device(clouds, z, S)
torch.cuda.synchronize()
acc = np.zeros(5)
N = 10
for it in range(N):
    flush.fill_(1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    ev[0].record()
    enc = pipe.encode(clouds); ev[1].record()
    rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(z, enc, 0.025, False); ev[2].record()
    p = m.DexEvaluator.native_forward(enc, rot, transl); ev[3].record()
    idx = torch.empty((B, 10), dtype=torch.int32, device="cuda"); cnt = torch.empty((B,), dtype=torch.int32, device="cuda")
    score = torch.empty((B, 10), dtype=torch.float32, device="cuda")
    _native.check(L.dexgg_filter_topk(p.data_ptr(), B * S, S, 0.5, 10, idx.data_ptr(), cnt.data_ptr(), None, score.data_ptr(), _native.stream_ptr(torch.device("cuda", 0))))
    ev[4].record()
    sel = idx.clamp_min(0).long().reshape(-1); valid = (idx >= 0).reshape(-1, 1).float()
    a = (rot.reshape(-1, 9).index_select(0, sel) * valid); b = transl.index_select(0, sel) * valid; c = joint.index_select(0, sel) * valid
    ev[5].record()
    torch.cuda.synchronize()
    acc += np.array([ev[i].elapsed_time(ev[i + 1]) for i in range(5)])
print("bps, gen, eva, filter, gather (ms):", (acc / N).round(3), "sum", (acc.sum() / N).round(3))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts = []
for it in range(N):
    flush.fill_(1)
    e0.record(); pipe.run_device(clouds, z, S); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
print("run_device ms:", np.round(ts, 3))
