"""GPU: the tcgen05 / TMEM / TMA fused linear kernel in isolation against torch (fp32 math on the bf16-rounded operands)."""
import pytest
import torch

from dex_gan_grasp_b200 import _native

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 32, 64), (300, 768, 560), (1000, 512, 256), (257, 21, 512), (4096, 768, 48),
                                   (130, 1536, 12288), (20000, 768, 560)])
def test_linear_bf16_vs_torch(M, N, K):
    L = _native.lib()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = torch.randn((M, K), device="cuda", generator=g).bfloat16()
    W = (torch.randn((N, K), device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias, shift = (torch.randn(N, device="cuda", generator=g) for _ in range(2))
    scale = 0.5 + torch.rand(N, device="cuda", generator=g)
    res = torch.randn((M, N), device="cuda", generator=g).bfloat16()
    ref = torch.nn.functional.leaky_relu((A.float() @ W.float().t() + bias) * scale + shift + res.float(), 0.2)
    ldc = (N + 7) // 8 * 8
    C16 = torch.zeros((M, ldc), device="cuda", dtype=torch.bfloat16)
    C32 = torch.zeros((M, ldc), device="cuda", dtype=torch.float32)
    _native.check(L.dexgg_linear_bf16(A.data_ptr(), K, W.data_ptr(), K, M, N, K, bias.data_ptr(), scale.data_ptr(), shift.data_ptr(),
                                      res.data_ptr(), N if N % 8 == 0 else 0, 1, C16.data_ptr(), C32.data_ptr(), ldc, None)
                  if N % 8 == 0 else
                  L.dexgg_linear_bf16(A.data_ptr(), K, W.data_ptr(), K, M, N, K, bias.data_ptr(), scale.data_ptr(), shift.data_ptr(),
                                      None, 0, 1, C16.data_ptr(), C32.data_ptr(), ldc, None))
    torch.cuda.synchronize()
    if N % 8 != 0:
        ref = torch.nn.functional.leaky_relu((A.float() @ W.float().t() + bias) * scale + shift, 0.2)
    err32 = (C32[:, :N] - ref).abs().max().item()
    assert err32 < 2e-3 * max(1.0, ref.abs().max().item()), err32   # fp32 accumulation order only
    err16 = (C16[:, :N].float() - ref).abs().max().item()
    assert err16 < 1e-2 * max(1.0, ref.abs().max().item()), err16   # + bf16 rounding of the output
    assert torch.all(C32[:, N:] == 0) and torch.all(C16[:, N:] == 0)  # nothing written outside [M,N]


@pytest.mark.parametrize("M,N,K", [(256, 768, 4101), (192, 256, 4613), (4096, 40, 517), (64, 4117, 768)])
def test_linear_bf16_any_k_with_padded_pitch(M, N, K):
    """K that is not a multiple of 8 (the first-layer widths 4101 / 4613 / 4117 of the training step): the row pitch is padded
    to 16 bytes and the pad columns hold garbage - the tensor map clips at K, so they must not reach the product."""
    L = _native.lib()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    Kp = (K + 7) // 8 * 8
    A = torch.full((M, Kp), float("nan"), device="cuda", dtype=torch.bfloat16)
    W = torch.full((N, Kp), float("nan"), device="cuda", dtype=torch.bfloat16)
    A[:, :K] = torch.randn((M, K), device="cuda", generator=g).bfloat16()
    W[:, :K] = (torch.randn((N, K), device="cuda", generator=g) / K ** 0.5).bfloat16()
    C32 = torch.zeros((M, N), device="cuda", dtype=torch.float32)
    _native.check(L.dexgg_linear_bf16(A.data_ptr(), Kp, W.data_ptr(), Kp, M, N, K, None, None, None, None, 0, 0, None, C32.data_ptr(), N, None))
    torch.cuda.synchronize()
    ref = A[:, :K].float() @ W[:, :K].float().t()
    assert torch.isfinite(C32).all()
    assert (C32 - ref).abs().max().item() < 2e-3 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("R,C", [(1, 1), (33, 65), (4096, 4101), (1000, 7)])
def test_transposing_convert(R, C):
    L = _native.lib()
    x = torch.randn((R, C + 3), device="cuda")[:, :C]  # strided source
    Rp = (R + 7) // 8 * 8
    out = torch.zeros((C, Rp), device="cuda", dtype=torch.bfloat16)
    _native.check(L.dexgg_f32_to_bf16_t(x.data_ptr(), x.stride(0), out.data_ptr(), Rp, R, C, None))
    torch.cuda.synchronize()
    assert torch.equal(out[:, :R], x.t().bfloat16())
    assert torch.all(out[:, R:] == 0)
