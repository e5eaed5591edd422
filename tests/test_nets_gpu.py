"""GPU parity of the generator / evaluator / discriminator forwards, rot6d, palm offset and the filter (all through the
C ABI) against (1) golden outputs of the reference's own nn.Modules (tests/golden/nets_golden.npz) and (2) the CPU oracle
(oracle/nets_oracle.py) on seeded inputs.  Tolerance for the fp32 mode: 1e-4 absolute (BASELINE.json north_star)."""
import copy

import numpy as np
import pytest
import torch

from dex_gan_grasp_b200 import _native, synth, tolerances
from oracle import nets_oracle as no

pytestmark = pytest.mark.gpu
TOL = tolerances.FP32_TOL
SEED = 7
CFG = dict(latentD=5, n_hand_joints=12, is_train=False, gpu_ids=[0], model="dexgangrasp", is_wgan=False)


def md(a, b):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    return float(np.max(np.abs(a.astype(np.float64).reshape(-1) - b.astype(np.float64).reshape(-1))))


@pytest.fixture(scope="module")
def model():
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    m.DexGANGrasp.eval()
    m.DexEvaluator.eval()
    return m


def test_linear_kernel_vs_torch():
    """The fused linear kernel in isolation, all operand layouts and the full epilogue, against torch fp32."""
    torch.backends.cuda.matmul.allow_tf32 = False
    L = _native.lib()
    g = torch.Generator(device="cuda").manual_seed(1)
    for (M, N, K, ta, tw) in [(300, 768, 528, 0, 0), (64, 21, 512, 0, 0), (257, 130, 4101, 0, 0), (128, 4101, 256, 0, 1),
                              (256, 4101, 300, 1, 1), (1, 1, 1, 0, 0), (1000, 512, 16, 0, 0)]:
        A = torch.randn((K, M) if ta else (M, K), device="cuda", generator=g)
        W = torch.randn((K, N) if tw else (N, K), device="cuda", generator=g) / K ** 0.5
        bias, scale, shift = (torch.randn(N, device="cuda", generator=g) for _ in range(3))
        res = torch.randn((M, N), device="cuda", generator=g)
        C = torch.randn((M, N), device="cuda", generator=g)
        C0 = C.clone()
        Am = A.t() if ta else A
        Wm = W.t() if tw else W
        ref = torch.nn.functional.leaky_relu((Am.double() @ Wm.double().t() + bias.double()) * scale.double() + shift.double()
                                             + res.double(), 0.2) + C0.double()
        _native.check(L.dexgg_linear_f32(A.data_ptr(), A.stride(0), ta, W.data_ptr(), W.stride(0), tw, M, N, K, bias.data_ptr(),
                                         scale.data_ptr(), shift.data_ptr(), res.data_ptr(), N, 1, 1, C.data_ptr(), N, None))
        torch.cuda.synchronize()
        assert md(C, ref) < 2e-5 * max(1.0, float(ref.abs().max())), (M, N, K, ta, tw)


def test_rowwise_forward_vs_reference_golden(model, eval_batch, nets_golden):
    """Config C2 shape: data/eval_batch.pth, every row its own BPS encoding (rows_per_obj = 1)."""
    bps = torch.from_numpy(eval_batch["bps_object"]).cuda()
    z = torch.from_numpy(nets_golden["A/z"]).cuda()
    out = model.DexGANGrasp.generator(z, bps)  # nn.Module surface: dict with rot_6D / transl / joint_conf / z
    assert set(out) == {"rot_6D", "transl", "joint_conf", "z"}
    assert md(out["rot_6D"], nets_golden["A/rot_6D"]) < TOL
    assert md(out["joint_conf"], nets_golden["A/joint_conf"]) < TOL
    rot, transl, joint, _ = model.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    assert md(rot, nets_golden["A/rot_matrix"]) < TOL and md(transl, nets_golden["A/transl"]) < TOL
    from dex_gan_grasp_b200 import utils
    assert md(utils.rot_matrix_from_ortho6d(out["rot_6D"]), nets_golden["A/rot_matrix"]) < TOL
    p = model.DexEvaluator({"bps_object": bps, "rot_matrix": torch.from_numpy(nets_golden["A/rot_matrix"]),
                            "transl": torch.from_numpy(nets_golden["A/transl"])})
    assert p.shape == (64, 1) and md(p, nets_golden["A/p_eva"]) < TOL
    real = {k: torch.from_numpy(eval_batch[k]) for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}  # f64 inputs
    assert md(model.DexGANGrasp.discriminator(real), nets_golden["A/p_disc_real"]) < TOL
    assert md(model.DexEvaluator(real), nets_golden["A/p_eva_real"]) < TOL
    fake = {"bps_object": bps, "rot_matrix": torch.from_numpy(nets_golden["A/rot_matrix"]),
            "transl": torch.from_numpy(nets_golden["A/transl"]), "joint_conf": torch.from_numpy(nets_golden["A/joint_conf"])}
    assert md(model.DexGANGrasp.discriminator(fake), nets_golden["A/p_disc_fake"]) < TOL


def test_shared_bps_hoisted_path_and_filter(model, nets_golden):
    """Config C1 shape: one object's encoding, n latent samples; the BPS projection is hoisted (once per object)."""
    z = torch.from_numpy(nets_golden["B/z"]).cuda()
    bps = torch.from_numpy(nets_golden["B/bps"]).cuda()  # [1,4096]
    rot, transl, joint, _ = model.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    assert md(rot, nets_golden["B/rot_matrix"]) < TOL and md(transl, nets_golden["B/transl"]) < TOL
    assert md(joint, nets_golden["B/joint_conf"]) < TOL
    grasps = {"rot_matrix": nets_golden["B/rot_matrix"].copy(), "transl": nets_golden["B/transl"].copy(),
              "joint_conf": nets_golden["B/joint_conf"].copy(), "z": nets_golden["B/z"].copy()}
    p = model.evaluate_grasps(nets_golden["B/bps"], dict(grasps))
    assert p.shape == (256,) and md(p, nets_golden["B/p_eva"]) < TOL
    thr = float(nets_golden["B/thresh_med"])
    f = model.filter_grasps(nets_golden["B/bps"], grasps, thresh=thr)
    assert "bps_object" in grasps and grasps["bps_object"].shape == (256, 4096)  # reference side effect (dexgangrasp.py:408)
    assert set(f) == {"rot_matrix", "transl", "joint_conf", "z", "bps_object"}
    # same kept rows in the same order as the reference, rows whose score is within TOL of the threshold excluded
    pg = nets_golden["B/p_eva"]
    sure = np.abs(pg - thr) > TOL
    kept_ref = [i for i in nets_golden["B/kept_med"] if sure[i]]
    kept = []
    for row in f["z"]:
        i = int(np.where((nets_golden["B/z"] == row).all(1))[0][0])
        if sure[i]:
            kept.append(i)
    gaps_ok = all(abs(pg[a] - pg[b]) < 2 * TOL for a, b in zip(kept, kept_ref) if a != b)  # near-equal scores may swap
    assert len(kept) == len(kept_ref) and gaps_ok
    assert f["bps_object"].shape == (len(f["z"]), 4096) and f["rot_matrix"].shape[1:] == (3, 3)
    with pytest.raises(ValueError):
        model.filter_grasps(nets_golden["B/bps"], dict(grasps), thresh=1.5)
    assert len(model.filter_grasps(nets_golden["B/bps"], dict(grasps), thresh=-1)["z"]) == 256  # pure sort (dexafford_prompt.py:198)


def test_wrapper_generate_matches_oracle_with_same_z(model, bps_golden):
    """generate_grasps draws z with torch.randn on the device exactly like networks.py:240; with that z the CPU oracle
    must give the same grasps; then filter_grasps_discriminator against the oracle discriminator."""
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    bps = bps_golden["mustard_bottle/bps"]
    torch.manual_seed(123)
    g = model.generate_grasps(bps, n_samples=100, return_arr=True)  # dexgangrasp_offline-style C1
    assert g["rot_matrix"].shape == (100, 3, 3) and g["transl"].dtype == np.float32 and g["z"].shape == (100, 5)
    torch.manual_seed(123)
    assert np.array_equal(torch.randn((100, 5), device="cuda").cpu().numpy(), g["z"])
    o = no.generate_poses(sd_gan, torch.from_numpy(g["z"]), torch.from_numpy(np.tile(bps.squeeze(), (100, 1))), 0.025)
    for k in ("rot_matrix", "transl", "joint_conf"):
        assert md(g[k], o[k]) < TOL, k
    with torch.no_grad():
        pd = no.discriminator_forward(sd_gan, torch.from_numpy(np.tile(bps.squeeze(), (100, 1))), torch.from_numpy(o["rot_matrix"]),
                                      torch.from_numpy(o["transl"]), torch.from_numpy(o["joint_conf"])).squeeze().numpy()
    thr = float(np.median(pd))
    f = model.filter_grasps_discriminator(bps, {k: g[k] for k in ("rot_matrix", "transl", "joint_conf", "z")}, thresh=thr)
    n_sure = int((pd > thr + TOL).sum())
    assert n_sure <= len(f["z"]) <= int((pd > thr - TOL).sum())


def test_many_objects_rows_per_object(model, bps_golden):
    """Batched form of config C3 at a size the oracle finishes in seconds: 7 objects x 300 samples in one call, plus the
    chunked path (more rows than one internal chunk) checked for self-consistency."""
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    names = list(bps_golden["names"])
    bps = np.concatenate([bps_golden[f"{n}/bps"] for n in names])  # [7,4096]
    S = 300
    z = synth.synth_latents(7 * S, 5, seed=21)
    rot, transl, joint, r6 = model.DexGANGrasp.generator.native_forward(z, torch.from_numpy(bps), 0.025, True)
    tiled = torch.from_numpy(np.repeat(bps, S, axis=0))
    o = no.generate_poses(sd_gan, z, tiled, 0.025)
    with torch.no_grad():
        r6o = no.generator_forward(sd_gan, z, tiled)["rot_6D"]
        po = no.evaluator_forward(sd_eva, tiled, torch.from_numpy(o["rot_matrix"]), torch.from_numpy(o["transl"]))
    p = model.DexEvaluator.native_forward(torch.from_numpy(bps), o["rot_matrix"], o["transl"])
    assert md(r6, r6o) < TOL and md(joint, o["joint_conf"]) < TOL and md(p, po) < TOL
    # Gram-Schmidt divides by |a| and |x × b|: the 1e-4 budget on the raw 6-D output is amplified by their inverse
    a, b = r6o[:, :3], r6o[:, 3:]
    x = a / a.norm(dim=1, keepdim=True)
    amp = (1.0 / torch.minimum(a.norm(dim=1), torch.linalg.cross(x, b).norm(dim=1))).clamp_min(1.0).numpy()
    drot = np.abs(rot.cpu().numpy() - o["rot_matrix"]).reshape(len(amp), -1).max(1)
    dtr = np.abs(transl.cpu().numpy() - o["transl"]).max(1)
    assert np.all(drot < 4 * TOL * amp) and np.all(dtr < TOL + 0.025 * 4 * TOL * amp)
    assert np.mean(drot < TOL) > 0.99
    # chunking: 2 objects x 40000 rows (> 32768-row chunk, chunk boundary inside an object) == same rows in small calls
    S2 = 40000
    z2 = synth.synth_latents(2 * S2, 5, seed=22)
    r_big, t_big, j_big, _ = model.DexGANGrasp.generator.native_forward(z2, torch.from_numpy(bps[:2]), 0.025, False)
    p_big = model.DexEvaluator.native_forward(torch.from_numpy(bps[:2]), r_big, t_big)
    sl = slice(S2 - 500, S2 + 500)  # straddles the object boundary
    for ob, lo, hi in ((0, S2 - 500, S2), (1, S2, S2 + 500)):
        r_s, t_s, j_s, _ = model.DexGANGrasp.generator.native_forward(z2[lo:hi], torch.from_numpy(bps[ob:ob + 1]), 0.025, False)
        p_s = model.DexEvaluator.native_forward(torch.from_numpy(bps[ob:ob + 1]), r_s, t_s)
        assert torch.equal(r_s, r_big[lo:hi]) and torch.equal(t_s, t_big[lo:hi]) and torch.equal(p_s, p_big[lo:hi])


def test_filter_topk_segments_vs_oracle():
    L = _native.lib()
    rng = np.random.default_rng(0)
    for (nseg, S, k, thr) in [(5, 1000, 10, 0.5), (3, 257, 257, 0.2), (1, 16384, 100, 0.9), (4, 64, 8, 2.0)]:
        p = rng.random(nseg * S).astype(np.float32)
        p[rng.integers(0, nseg * S, nseg * S // 4)] = np.float32(0.75)  # many exact ties
        pt = torch.from_numpy(p).cuda()
        idx = torch.empty((nseg, k), dtype=torch.int32, device="cuda")
        cnt = torch.empty((nseg,), dtype=torch.int32, device="cuda")
        best = torch.empty((nseg,), dtype=torch.float32, device="cuda")
        sc = torch.empty((nseg, k), dtype=torch.float32, device="cuda")
        _native.check(L.dexgg_filter_topk(pt.data_ptr(), nseg * S, S, thr, k, idx.data_ptr(), cnt.data_ptr(), best.data_ptr(),
                                          sc.data_ptr(), None))
        io, co = no.segmented_topk(p, np.arange(nseg + 1) * S, thr, k)
        assert np.array_equal(idx.cpu().numpy(), io) and np.array_equal(cnt.cpu().numpy(), co)
        assert np.array_equal(best.cpu().numpy(), p.reshape(nseg, S).max(1))
        m = io >= 0
        assert np.array_equal(sc.cpu().numpy()[m], p[io[m]])


def test_filter_topk_large_segments_and_gather():
    """segments longer than the 16384-row shared-memory sort (eval.py samples 5 x n_gt grasps per object) go
    through the workspace path; and the one-launch top-k + gather of the batched pipeline returns exact int32 row ids."""
    L = _native.lib()
    rng = np.random.default_rng(1)
    for (nseg, S, k, thr) in [(1, 16385, 16385, 0.3), (1, 50000, 50000, 0.5), (2, 40000, 77, 0.9), (1, 100000, 100000, -1.0)]:
        p = rng.random(nseg * S).astype(np.float32)
        p[rng.integers(0, nseg * S, nseg * S // 8)] = np.float32(0.625)  # exact ties
        pt = torch.from_numpy(p).cuda()
        idx = torch.empty((nseg, k), dtype=torch.int32, device="cuda")
        cnt = torch.empty((nseg,), dtype=torch.int32, device="cuda")
        best = torch.empty((nseg,), dtype=torch.float32, device="cuda")
        nws = L.dexgg_filter_topk_workspace_bytes(nseg * S, S)
        assert nws > 0
        ws = torch.empty((nws,), dtype=torch.uint8, device="cuda")
        rc = L.dexgg_filter_topk(pt.data_ptr(), nseg * S, S, thr, k, idx.data_ptr(), cnt.data_ptr(), best.data_ptr(), None, None)
        assert rc == -3  # DEXGG_E_WORKSPACE: the plain entry point refuses long segments loudly
        _native.check(L.dexgg_filter_topk_ws(pt.data_ptr(), nseg * S, S, thr, k, idx.data_ptr(), cnt.data_ptr(), best.data_ptr(), None,
                                             ws.data_ptr(), nws, None))
        io, co = no.segmented_topk(p, np.arange(nseg + 1) * S, thr, k)
        assert np.array_equal(cnt.cpu().numpy(), co) and np.array_equal(idx.cpu().numpy(), io), (nseg, S)
        assert np.array_equal(best.cpu().numpy(), p.reshape(nseg, S).max(1))
    # gather: [n_seg, k, 14 + nJ] = score | row id bits | rot | transl | joint
    nseg, S, k, nJ = 37, 1000, 10, 12
    M = nseg * S
    p = rng.random(M).astype(np.float32)
    p[3 * S:4 * S] = 0.1  # one object with nothing above the threshold
    rot, transl, joint = (rng.standard_normal((M, c)).astype(np.float32) for c in (9, 3, nJ))
    d = [torch.from_numpy(a).cuda() for a in (p, rot, transl, joint)]
    idx = torch.empty((nseg, k), dtype=torch.int32, device="cuda")
    cnt = torch.empty((nseg,), dtype=torch.int32, device="cuda")
    packed = torch.empty((nseg, k, 14 + nJ), dtype=torch.float32, device="cuda")
    _native.check(L.dexgg_filter_topk_gather(d[0].data_ptr(), M, S, 0.5, k, d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), nJ,
                                             idx.data_ptr(), cnt.data_ptr(), packed.data_ptr(), None))
    io, co = no.segmented_topk(p, np.arange(nseg + 1) * S, 0.5, k)
    assert np.array_equal(idx.cpu().numpy(), io) and np.array_equal(cnt.cpu().numpy(), co) and co[3] == 0
    from dex_gan_grasp_b200.pipeline import unpack_host
    u = unpack_host(packed.cpu())
    assert np.array_equal(u["idx"].numpy(), io)
    m = io >= 0
    assert np.array_equal(u["score"].numpy()[m], p[io[m]]) and np.array_equal(u["rot_matrix"].numpy().reshape(nseg, k, 9)[m], rot[io[m]])
    assert np.array_equal(u["transl"].numpy()[m], transl[io[m]]) and np.array_equal(u["joint_conf"].numpy()[m], joint[io[m]])
    assert not u["rot_matrix"].numpy().reshape(nseg, k, 9)[~m].any() and not u["score"].numpy()[~m].any()


def test_weights_follow_parameter_updates(model, eval_batch, nets_golden):
    """The packed copies are re-derived when parameters change in place (optimizer step / load_state_dict)."""
    bps = torch.from_numpy(eval_batch["bps_object"]).cuda()
    z = torch.from_numpy(nets_golden["A/z"]).cuda()
    before = model.DexGANGrasp.generator(z, bps)["transl"].clone()
    with torch.no_grad():
        model.DexGANGrasp.generator.gen_transl.bias.add_(1.0)
    after = model.DexGANGrasp.generator(z, bps)["transl"]
    assert md(after, before + 1.0) < 1e-6
    model.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    assert md(model.DexGANGrasp.generator(z, bps)["transl"], before) == 0.0


def test_checkpoint_layout_roundtrip(model, tmp_path):
    """Checkpoint dict keys and state-dict keys are the reference's."""
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    cfg = dict(CFG, is_train=True, bce_weight=1.0, kl_coef=0.0, lr=1e-4, lr_gen=1e-4, lr_dis=1e-4, beta1=0.9, weight_decay=5e-4,
               lr_policy="plateau", threshold_lr_policy_plateau=0.01, patience_lr_policy_plateau=3, weight_init_type="xavier",
               init_gain=0.02, train_ffhgenerator=True, train_ffhevaluator=True, save_dir=str(tmp_path), load_path=str(tmp_path))
    m = DexGanGrasp(cfg)
    assert list(m.DexGANGrasp.state_dict().keys()) == list(synth.synth_gan_state_dict(SEED).keys())
    assert list(m.DexEvaluator.state_dict().keys()) == list(synth.synth_eva_state_dict(SEED).keys())
    m.save_dexgenerator("3", 3)
    m.save_dexevaluator("3", 3)
    ck = torch.load(tmp_path / "3_gen_net.pt")
    assert set(ck) == {"epoch", "ffhgenerator_state_dict", "optim_ffhgenerator_state_dict", "scheduler_ffhgenerator_state_dict",
                       "optim_ffhgan_generator_state_dict", "scheduler_ffhgan_generator_state_dict",
                       "optim_ffhgan_discriminator_state_dict", "scheduler_ffhgan_discriminator_state_dict"}
    assert set(torch.load(tmp_path / "3_eva_net.pt")) == {"epoch", "ffhevaluator_state_dict", "optim_ffhevaluator_state_dict",
                                                         "scheduler_ffhevaluator_state_dict"}
    m2 = DexGanGrasp(dict(cfg))
    m2.load_dexgenerator(3)
    m2.load_dexevaluator(3)
    for a, b in zip(m.DexGANGrasp.state_dict().values(), m2.DexGANGrasp.state_dict().values()):
        assert torch.equal(a, b)


# ---------------------------------------------------------------------------------------------------------------------
# bf16 tensor-core mode (tcgen05): hidden activations and weights are bf16, accumulation fp32; BPS / z / pose inputs are
# carried as hi+lo bf16 pairs.  Stated tolerance (BASELINE.json north_star "bf16 mode ... within a stated score tolerance"):
#   evaluator / discriminator scores within BF16_SCORE_TOL of the reference fp32 scores, generator outputs within
#   BF16_POSE_TOL, and the filtered grasp set identical except rows whose fp32 score lies within BF16_SCORE_TOL of thresh.
BF16_SCORE_TOL = tolerances.BF16_SCORE_TOL  # one stage on identical inputs; the pipeline-level epsilon is
BF16_POSE_TOL = tolerances.BF16_POSE_TOL    # tolerances.BF16_PIPELINE_EPS (tests/test_pipeline_gpu.py)


@pytest.fixture(scope="module")
def model_bf16():
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG, precision="bf16"))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    m.DexGANGrasp.eval()
    m.DexEvaluator.eval()
    return m


def test_bf16_mode_rowwise_vs_reference_golden(model_bf16, eval_batch, nets_golden):
    m = model_bf16
    bps = torch.from_numpy(eval_batch["bps_object"]).cuda()
    z = torch.from_numpy(nets_golden["A/z"]).cuda()
    rot, transl, joint, r6 = m.DexGANGrasp.generator.native_forward(z, bps, 0.025, True)
    d = {"rot_6D": md(r6, nets_golden["A/rot_6D"]), "transl": md(transl, nets_golden["A/transl"]),
         "joint": md(joint, nets_golden["A/joint_conf"])}
    p = m.DexEvaluator({"bps_object": bps, "rot_matrix": torch.from_numpy(nets_golden["A/rot_matrix"]),
                        "transl": torch.from_numpy(nets_golden["A/transl"])})
    real = {k: torch.from_numpy(eval_batch[k]) for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}
    d["p_eva"] = md(p, nets_golden["A/p_eva"])
    d["p_disc_real"] = md(m.DexGANGrasp.discriminator(real), nets_golden["A/p_disc_real"])
    print("bf16 vs reference fp32 (row-wise):", d)
    assert d["rot_6D"] < BF16_POSE_TOL and d["transl"] < BF16_POSE_TOL and d["joint"] < BF16_POSE_TOL
    assert d["p_eva"] < BF16_SCORE_TOL and d["p_disc_real"] < BF16_SCORE_TOL


def test_bf16_mode_same_filtered_set(model_bf16, nets_golden):
    m = model_bf16
    z = torch.from_numpy(nets_golden["B/z"]).cuda()
    bps = torch.from_numpy(nets_golden["B/bps"]).cuda()
    rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    assert md(transl, nets_golden["B/transl"]) < BF16_POSE_TOL and md(joint, nets_golden["B/joint_conf"]) < BF16_POSE_TOL
    grasps = {"rot_matrix": nets_golden["B/rot_matrix"].copy(), "transl": nets_golden["B/transl"].copy(),
              "joint_conf": nets_golden["B/joint_conf"].copy(), "z": nets_golden["B/z"].copy()}
    p = m.evaluate_grasps(nets_golden["B/bps"], dict(grasps))
    pg = nets_golden["B/p_eva"]
    print("bf16 scores vs reference fp32 (shared BPS): max |dp| =", float(np.max(np.abs(p - pg))))
    assert np.max(np.abs(p - pg)) < BF16_SCORE_TOL
    thr = float(nets_golden["B/thresh_med"])
    f = m.filter_grasps(nets_golden["B/bps"], grasps, thresh=thr)
    kept = {int(np.where((nets_golden["B/z"] == row).all(1))[0][0]) for row in f["z"]}
    sure_in = {i for i in range(len(pg)) if pg[i] > thr + BF16_SCORE_TOL}
    sure_out = {i for i in range(len(pg)) if pg[i] < thr - BF16_SCORE_TOL}
    assert sure_in <= kept and not (kept & sure_out)


def test_bf16_mode_many_objects_vs_fp32_mode(model, model_bf16, bps_golden):
    names = list(bps_golden["names"])
    bps = torch.from_numpy(np.concatenate([bps_golden[f"{n}/bps"] for n in names]))
    S = 3000  # 21000 rows: more than one 148x128-row chunk
    z = synth.synth_latents(7 * S, 5, seed=31)
    out32 = model.DexGANGrasp.generator.native_forward(z, bps, 0.025, True)
    out16 = model_bf16.DexGANGrasp.generator.native_forward(z, bps, 0.025, True)
    assert md(out16[3], out32[3]) < BF16_POSE_TOL and md(out16[1], out32[1]) < BF16_POSE_TOL and md(out16[2], out32[2]) < BF16_POSE_TOL
    p32 = model.DexEvaluator.native_forward(bps, out32[0], out32[1])
    p16 = model_bf16.DexEvaluator.native_forward(bps, out32[0], out32[1])
    d32 = model.DexGANGrasp.discriminator.native_forward(bps, out32[0], out32[1], out32[2])
    d16 = model_bf16.DexGANGrasp.discriminator.native_forward(bps, out32[0], out32[1], out32[2])
    print("bf16 vs fp32 mode: max |dp_eva| =", md(p16, p32), " max |dp_disc| =", md(d16, d32), " r6", md(out16[3], out32[3]))
    assert md(p16, p32) < BF16_SCORE_TOL and md(d16, d32) < BF16_SCORE_TOL


def test_bf16_discriminator_runs_on_the_fused_chain_kernel(model, model_bf16, bps_golden):
    """The bf16 discriminator forward (filter_grasps_discriminator's scorer, networks.py:148-159) goes through the fused whole-network
    kernel like the generator and the evaluator: a handful of launches (projection GEMM, base finish, chain kernel) instead of one
    per layer, ragged M (not a multiple of 128, tiles over two objects), scores within BF16_SCORE_TOL of the fp32 mode."""
    from dex_gan_grasp_b200 import _native
    names = list(bps_golden["names"])
    bps = torch.from_numpy(np.concatenate([bps_golden[f"{n}/bps"] for n in names]))
    S = 333
    z = synth.synth_latents(7 * S, 5, seed=5)
    rot, transl, joint, _ = model.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    d32 = model.DexGANGrasp.discriminator.native_forward(bps, rot, transl, joint)
    model_bf16.DexGANGrasp.discriminator.native_forward(bps, rot, transl, joint)  # packs the weights
    torch.cuda.synchronize()
    n0 = _native.lib().dexgg_launch_count()
    d16 = model_bf16.DexGANGrasp.discriminator.native_forward(bps, rot, transl, joint)
    torch.cuda.synchronize()
    launches = _native.lib().dexgg_launch_count() - n0
    assert launches <= 8, launches
    assert md(d16, d32) < BF16_SCORE_TOL


def test_pipeline_streaming_equals_one_step_at_a_time(model):
    """GraspPipeline.run_host_pipelined (copies of neighbouring steps overlap the compute, rotating buffers) returns, step by
    step, exactly what run_host returns for the same inputs."""
    from dex_gan_grasp_b200.pipeline import GraspPipeline
    pipe = GraspPipeline(model.DexGANGrasp, model.DexEvaluator, k=5, thresh=0.5, z_offset=0.025)
    B, S = 6, 200
    steps = []
    for i in range(5):
        clouds = torch.from_numpy(synth.synth_clouds(B, 1024, seed0=50 + 10 * i)).pin_memory()
        z = torch.randn((B * S, 5), generator=torch.Generator().manual_seed(i)).pin_memory()
        steps.append((clouds, z))
    want = [pipe.run_host(c, z, S)[0] for c, z in steps]
    want = [{k: v.clone() for k, v in w.items()} for w in want]
    n = 0
    for (res, h2d, d2h), w in zip(pipe.run_host_pipelined(iter(steps), S), want):
        assert torch.equal(res["count"], w["count"]) and torch.equal(res["packed"], w["packed"]), n
        assert h2d == steps[0][0].numel() * 4 + steps[0][1].numel() * 4 and d2h > 0
        n += 1
    assert n == len(steps)


# ---------------------------------------------------------------------------------------------------------------------
# fp32_tc: the 1e-4 contract ON the tensor cores (every activation / weight as a bf16 hi + lo pair, hi*hi + lo*hi + hi*lo per
# MAC in one tcgen05 GEMM, fp32 skip branch).  Same golden vectors and the same TOL as the CUDA-core fp32 mode.
@pytest.fixture(scope="module")
def model_tc():
    from dex_gan_grasp_b200.dexgangrasp import DexGanGrasp
    m = DexGanGrasp(dict(CFG, precision="fp32_tc"))
    m.DexGANGrasp.load_state_dict(synth.synth_gan_state_dict(SEED))
    m.DexEvaluator.load_state_dict(synth.synth_eva_state_dict(SEED))
    m.DexGANGrasp.eval()
    m.DexEvaluator.eval()
    return m


def test_fp32_tc_rowwise_vs_reference_golden(model_tc, eval_batch, nets_golden):
    m = model_tc
    bps = torch.from_numpy(eval_batch["bps_object"]).cuda()
    z = torch.from_numpy(nets_golden["A/z"]).cuda()
    rot, transl, joint, r6 = m.DexGANGrasp.generator.native_forward(z, bps, 0.025, True)
    d = {"rot_6D": md(r6, nets_golden["A/rot_6D"]), "transl": md(transl, nets_golden["A/transl"]), "joint": md(joint, nets_golden["A/joint_conf"])}
    # Gram-Schmidt divides by |a| and |x x b|: the 1e-4 budget on the raw 6-D output is amplified by their inverse (as in
    # test_many_objects_rows_per_object for the CUDA-core mode)
    r6g = torch.from_numpy(nets_golden["A/rot_6D"])
    a, b = r6g[:, :3], r6g[:, 3:]
    x = a / a.norm(dim=1, keepdim=True)
    amp = (1.0 / torch.minimum(a.norm(dim=1), torch.linalg.cross(x, b).norm(dim=1))).clamp_min(1.0).numpy()
    drot = np.abs(rot.cpu().numpy() - nets_golden["A/rot_matrix"]).reshape(len(amp), -1).max(1)
    assert np.all(drot < 4 * TOL * amp), float((drot / amp).max())
    p = m.DexEvaluator({"bps_object": bps, "rot_matrix": torch.from_numpy(nets_golden["A/rot_matrix"]),
                        "transl": torch.from_numpy(nets_golden["A/transl"])})
    real = {k: torch.from_numpy(eval_batch[k]) for k in ("bps_object", "rot_matrix", "transl", "joint_conf")}
    d["p_eva"] = md(p, nets_golden["A/p_eva"])
    d["p_disc_real"] = md(m.DexGANGrasp.discriminator(real), nets_golden["A/p_disc_real"])
    d["p_eva_real"] = md(m.DexEvaluator(real), nets_golden["A/p_eva_real"])
    print("fp32_tc vs reference fp32 (row-wise):", d)
    assert all(v < TOL for v in d.values()), d


def test_fp32_tc_shared_bps_and_many_objects(model_tc, nets_golden, bps_golden):
    m = model_tc
    z = torch.from_numpy(nets_golden["B/z"]).cuda()
    bps = torch.from_numpy(nets_golden["B/bps"]).cuda()
    rot, transl, joint, _ = m.DexGANGrasp.generator.native_forward(z, bps, 0.025, False)
    assert md(transl, nets_golden["B/transl"]) < TOL and md(joint, nets_golden["B/joint_conf"]) < TOL
    rot_o = torch.from_numpy(nets_golden["B/rot_matrix"])  # golden has no raw rot6d for case B: x = R[:,:,0] is a / |a|, so bound the
    # error by the same amplification estimated from the oracle's own 6-D output on these inputs
    sd_gan0 = synth.synth_gan_state_dict(SEED)
    with torch.no_grad():
        r6b = no.generator_forward(sd_gan0, torch.from_numpy(nets_golden["B/z"]), torch.from_numpy(np.tile(nets_golden["B/bps"], (256, 1))))["rot_6D"]
    a, b = r6b[:, :3], r6b[:, 3:]
    x = a / a.norm(dim=1, keepdim=True)
    amp = (1.0 / torch.minimum(a.norm(dim=1), torch.linalg.cross(x, b).norm(dim=1))).clamp_min(1.0).numpy()
    drot = np.abs(rot.cpu().numpy() - rot_o.numpy()).reshape(len(amp), -1).max(1)
    assert np.all(drot < 4 * TOL * amp), float((drot / amp).max())
    grasps = {"rot_matrix": nets_golden["B/rot_matrix"].copy(), "transl": nets_golden["B/transl"].copy(),
              "joint_conf": nets_golden["B/joint_conf"].copy(), "z": nets_golden["B/z"].copy()}
    p = m.evaluate_grasps(nets_golden["B/bps"], dict(grasps))
    assert md(p, nets_golden["B/p_eva"]) < TOL
    # 7 objects x 3000 samples (more than one 148 x 128-row chunk) against the oracle
    sd_gan, sd_eva = synth.synth_gan_state_dict(SEED), synth.synth_eva_state_dict(SEED)
    names = list(bps_golden["names"])
    b7 = np.concatenate([bps_golden[f"{n}/bps"] for n in names])
    S = 3000
    zz = synth.synth_latents(7 * S, 5, seed=41)
    rot, transl, joint, r6 = m.DexGANGrasp.generator.native_forward(zz, torch.from_numpy(b7), 0.025, True)
    tiled = torch.from_numpy(np.repeat(b7, S, axis=0))
    o = no.generate_poses(sd_gan, zz, tiled, 0.025)
    with torch.no_grad():
        r6o = no.generator_forward(sd_gan, zz, tiled)["rot_6D"]
        po = no.evaluator_forward(sd_eva, tiled, torch.from_numpy(o["rot_matrix"]), torch.from_numpy(o["transl"]))
    pe = m.DexEvaluator.native_forward(torch.from_numpy(b7), o["rot_matrix"], o["transl"])
    pd = m.DexGANGrasp.discriminator.native_forward(torch.from_numpy(b7), o["rot_matrix"], o["transl"], o["joint_conf"])
    with torch.no_grad():
        pdo = no.discriminator_forward(sd_gan, tiled, torch.from_numpy(o["rot_matrix"]), torch.from_numpy(o["transl"]), torch.from_numpy(o["joint_conf"]))
    d = {"r6": md(r6, r6o), "joint": md(joint, o["joint_conf"]), "p_eva": md(pe, po), "p_disc": md(pd, pdo)}
    print("fp32_tc vs oracle (7 x 3000):", d)
    assert all(v < TOL for v in d.values()), d
