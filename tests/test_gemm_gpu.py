"""dexgg_gemm_bf16 (csrc/gemm_umma.cu: the training step's general tcgen05 GEMM) against torch on the same bf16 operands,
fp32 accumulation: every operand-layout combination (K-major / MN-major for A and B), K segments, ragged M / N / K, every
N tile, bias + LeakyReLU column range, bf16 / fp32 outputs, accumulate and the split contraction (atomic adds)."""
import ctypes

import pytest
import torch

from dex_gan_grasp_b200 import _native

pytestmark = pytest.mark.gpu


def run_gemm(segs, M, N, bias=None, act_from=None, out16=False, out32=True, accumulate=False, split_k=0, tile_n=0, c32=None):
    """segs: list of (A_tensor, a_mn, B_tensor, b_mn, K) with bf16 tensors in their storage layout."""
    g = _native.Gemm()
    g.M, g.N, g.n_seg = M, N, len(segs)
    for i, (A, a_mn, B, b_mn, K) in enumerate(segs):
        g.seg[i].A, g.seg[i].lda, g.seg[i].a_mn_major = A.data_ptr(), A.stride(0), int(a_mn)
        g.seg[i].B, g.seg[i].ldb, g.seg[i].b_mn_major = B.data_ptr(), B.stride(0), int(b_mn)
        g.seg[i].K = K
    g.bias = bias.data_ptr() if bias is not None else None
    g.act_from = N if act_from is None else act_from
    ldc = (N + 7) // 8 * 8
    C16 = torch.zeros((M, ldc), dtype=torch.bfloat16, device="cuda") if out16 else None
    C32 = (c32 if c32 is not None else torch.zeros((M, ldc), dtype=torch.float32, device="cuda")) if out32 else None
    g.C16 = C16.data_ptr() if out16 else None
    g.C32 = C32.data_ptr() if out32 else None
    g.ldc, g.accumulate, g.split_k, g.tile_n = ldc, int(accumulate), split_k, tile_n
    _native.check(_native.lib().dexgg_gemm_bf16(ctypes.byref(g), None))
    torch.cuda.synchronize()
    return (C16[:, :N] if out16 else None), (C32[:, :N] if out32 else None)


def make_operand(rows_i, K, mn_major, gen, scale=1.0, pad=0):
    """logical [rows_i, K] operand; storage [rows_i, K(+pad)] (K-major) or [K, rows_i(+pad)] (MN-major); returns (storage, logical fp32)."""
    logical = (torch.randn((rows_i, K), device="cuda", generator=gen) * scale).to(torch.bfloat16)
    if mn_major:
        ld = (rows_i + pad + 7) // 8 * 8
        st = torch.zeros((K, ld), dtype=torch.bfloat16, device="cuda")
        st[:, :rows_i] = logical.t()
    else:
        ld = (K + pad + 7) // 8 * 8
        st = torch.zeros((rows_i, ld), dtype=torch.bfloat16, device="cuda")
        st[:, :K] = logical
    return st, logical.float()


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_operand_layouts_vs_torch(a_mn, b_mn):
    gen = torch.Generator(device="cuda").manual_seed(10 * a_mn + b_mn)
    for (M, N, K) in [(128, 64, 64), (256, 192, 128), (300, 768, 520), (4096, 256, 96), (77, 32, 4101), (512, 544, 768), (32, 512, 1000),
                      (768, 4096, 300), (129, 130, 16)]:
        if b_mn and N < 64:
            continue  # MN-major B needs 64-column tiles (the heads use K-major B or N padded to 64)
        A, Al = make_operand(M, K, a_mn, gen, pad=8)
        B, Bl = make_operand(N, K, b_mn, gen, scale=K ** -0.5, pad=16)
        _, C = run_gemm([(A, a_mn, B, b_mn, K)], M, N)
        ref = Al.double() @ Bl.double().t()
        err = float((C.double() - ref).abs().max())
        assert err < 2e-3 * max(1.0, float(ref.abs().max())), (a_mn, b_mn, M, N, K, err)


def test_segments_bias_act_outputs_and_tiles():
    gen = torch.Generator(device="cuda").manual_seed(5)
    M, N = 1000, 768
    Ks = [4096, 544]
    segs, ref = [], torch.zeros((M, N), dtype=torch.float64, device="cuda")
    for K in Ks:
        A, Al = make_operand(M, K, 0, gen)
        B, Bl = make_operand(N, K, 0, gen, scale=K ** -0.5)
        segs.append((A, 0, B, 0, K))
        ref += Al.double() @ Bl.double().t()
    bias = torch.randn(N, device="cuda", generator=gen)
    ref = ref + bias.double()
    ref_act = ref.clone()
    ref_act[:, 256:] = torch.nn.functional.leaky_relu(ref[:, 256:], 0.2)
    for tile_n in (0, 128, 192, 256):
        C16, C32 = run_gemm(segs, M, N, bias=bias, act_from=256, out16=True, out32=True, tile_n=tile_n)
        assert float((C32.double() - ref_act).abs().max()) < 3e-3 * float(ref.abs().max()), tile_n
        assert float((C16.double() - ref_act).abs().max()) < 1e-2 * float(ref.abs().max()), tile_n
    # accumulate into an existing fp32 buffer
    base = torch.randn((M, N), device="cuda", generator=gen)
    c32 = base.clone()
    _, C = run_gemm(segs, M, N, accumulate=True, c32=c32)
    assert float((C.double() - (ref - bias.double() + base.double())).abs().max()) < 3e-3 * float(ref.abs().max())


def test_weight_gradient_shapes_with_split_contraction():
    """dW = dY^T . X with both operands read MN-major from their row-major [rows, features] storage; the small layers split the
    contraction over CTAs (atomic accumulation)."""
    gen = torch.Generator(device="cuda").manual_seed(6)
    for (rows, n_out, k_in, split) in [(4096, 512, 256, 16), (8192, 512, 256, 32), (4096, 32, 512, 8), (1000, 768, 544, 1), (4096, 768, 4096, 1),
                                       (192, 512, 256, 3)]:
        dY = (torch.randn((rows, n_out), device="cuda", generator=gen) * rows ** -0.5).to(torch.bfloat16)
        X = torch.randn((rows, k_in), device="cuda", generator=gen).to(torch.bfloat16)
        _, dW = run_gemm([(dY, 1, X, 1, rows)], n_out, k_in, split_k=split)
        ref = dY.double().t() @ X.double()
        err = float((dW.double() - ref).abs().max())
        assert err < 2e-3 * max(1.0, float(ref.abs().max())), (rows, n_out, k_in, split, err)
    # dgrad: dX = dY . W with W [n_out, k_in] read as the MN-major B operand
    dY = torch.randn((4096, 768), device="cuda", generator=gen).to(torch.bfloat16)
    W = (torch.randn((768, 544), device="cuda", generator=gen) * 768 ** -0.5).to(torch.bfloat16)
    _, dX = run_gemm([(dY, 0, W, 1, 768)], 4096, 544)
    ref = dY.double() @ W.double()
    assert float((dX.double() - ref).abs().max()) < 2e-3 * float(ref.abs().max())
