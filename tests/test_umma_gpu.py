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
