"""Measured error of the tensor-core form of a(k) = |x_k|^2 - 2 b.x_k used by the BPS filter scan (csrc/bps_umma.cu): the same
three-way bf16 split of the coordinates, the same K = 32 slot order and the same tcgen05 instruction (kind::f16, fp32 accumulate,
through dexgg_gemm_bf16) against float64, in units of u (3 C^2 + 2 |b|_1 C) with u = 2^-24 - the kernel's margin assumes <= 72."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dex_gan_grasp_b200 import bps as bps_mod, synth
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gemm_gpu import run_gemm


def split3(v):  # fp32 tensor -> three bf16 tensors (as fp32 values)
    p0 = v.to(torch.bfloat16).float()
    r1 = v - p0
    p1 = r1.to(torch.bfloat16).float()
    r2 = r1 - p1
    p2 = r2.to(torch.bfloat16).float()
    return p0, p1, p2


def operands(basis, pts):
    P, N = basis.shape[0], pts.shape[0]
    A = torch.zeros((P, 32), device="cuda")
    B = torch.zeros((N, 32), device="cuda")
    for c in range(3):
        bp = split3(basis[:, c])
        tp = split3(-2.0 * pts[:, c])
        for i in range(3):
            for j in range(3):
                A[:, 9 * c + 3 * i + j] = bp[j]
                B[:, 9 * c + 3 * i + j] = tp[i]
    x, y, z = pts[:, 0], pts[:, 1], pts[:, 2]
    w = torch.addcmul(torch.addcmul(x * x, y, y), z, z)  # fma(z,z,fma(y,y,x*x)) up to fusion; compared against THIS w below
    n = split3(w)
    for i in range(3):
        A[:, 27 + i] = 1.0
        B[:, 27 + i] = n[i]
    return A.to(torch.bfloat16), B.to(torch.bfloat16), w


basis = torch.from_numpy(np.load(bps_mod.BASIS_PATH)).cuda()
out = {}
u = 2.0 ** -24
cases = {"synth surfaces": torch.from_numpy(synth.synth_clouds(4, 2048).reshape(-1, 3)).cuda(),
         "gaussian 0.08": 0.08 * torch.randn((8192, 3), device="cuda"),
         "off-centre (4,-7,2.5)": 0.05 * torch.randn((8192, 3), device="cuda") + torch.tensor([4.0, -7.0, 2.5], device="cuda"),
         "huge sigma 30": 30.0 * torch.randn((8192, 3), device="cuda")}
for name, pts in cases.items():
    pts = pts.float().contiguous()
    A, B, w = operands(basis, pts)
    _, C = run_gemm([(A, 0, B, 0, 32)], basis.shape[0], pts.shape[0])
    exact = w.double()[None, :] - 2.0 * (basis.double() @ pts.double().t())
    C_ = float(pts.abs().max())
    b1 = float(basis.abs().sum(1).max())
    scale = u * (3 * C_ * C_ + 2 * b1 * C_)
    err = (C.double() - exact).abs()
    out[name] = {"max_err_in_u_units": float(err.max()) / scale, "mean_err_in_u_units": float(err.mean()) / scale, "C": C_, "b1": b1}
print(json.dumps(out, indent=1))
