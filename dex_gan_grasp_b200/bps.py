"""Host-side mirror of the reference's BPS interface, running on the sm_100a kernel of csrc/bps_encode.cu.

Mirrors
  * `bps_torch.bps.bps_torch(custom_basis=..., device=...).encode(x, feature_type=['dists'])` as the reference calls it
    (DexGanGrasp/data/bps_encoder.py:19,33; dexgangrasp_real_robot.py:101,139; dexafford_prompt.py:123,177), and
  * `DexGanGrasp.data.bps_encoder.BPSEncoder` (bps_encoder.py:12-55).
"""
import os

import numpy as np
import torch

from . import _native

BASIS_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "basis_point_set.npy")


def to_np(t):
    """bps_torch.utils.to_np"""
    if torch.is_tensor(t):
        return t.detach().cpu().numpy()
    return np.asarray(t)


class _Workspace:
    """Grow-only device scratch buffer (one per encoder; calls are stream-ordered on the current stream)."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes, device):
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        return self.buf


def bps_encode_device(basis, pts, offsets=None, n_clouds=None, max_cloud_points=None, return_idx=True, flags=0, ws=None):
    """basis [P,3] f32 cuda; pts [total,3] f32 cuda (clouds back to back) or [B,N,3]; offsets [B+1] i64 cuda or None
    (uniform).  Returns (dists [B,P] f32, idx [B,P] i32 or None) on the device; stream-ordered, no sync."""
    L = _native.lib()
    assert basis.is_cuda and pts.is_cuda and basis.dtype == torch.float32 and pts.dtype == torch.float32
    basis = basis.contiguous()
    pts = pts.contiguous()
    P = basis.shape[0]
    if offsets is None:
        assert pts.dim() == 3, "uniform clouds: pts must be [B,N,3]"
        B, maxn = pts.shape[0], pts.shape[1]
        total = B * maxn
    else:
        B = offsets.numel() - 1 if n_clouds is None else n_clouds
        total = pts.numel() // 3
        assert max_cloud_points is not None, "ragged clouds: pass max_cloud_points (host int)"
        maxn = int(max_cloud_points)
        assert offsets.is_cuda and offsets.dtype == torch.int64
    dists = torch.empty((B, P), dtype=torch.float32, device=pts.device)
    idx = torch.empty((B, P), dtype=torch.int32, device=pts.device) if return_idx else None
    nbytes = L.dexgg_bps_workspace_bytes(B, P)
    wbuf = (ws or _Workspace()).get(nbytes, pts.device)
    with torch.cuda.device(pts.device):
        _native.check(L.dexgg_bps_encode(_native.ptr(basis), P, _native.ptr(pts), _native.ptr(offsets), B, maxn, total,
                                         _native.ptr(dists), _native.ptr(idx), flags, _native.ptr(wbuf), wbuf.numel(),
                                         _native.stream_ptr(pts.device)))
    return dists, idx


def center_clouds_device(pts, offsets=None, max_cloud_points=None, out=None):
    """obj_pcd.translate(-obj_pcd.get_center()) (dexgangrasp_real_robot.py:133-134) on the device: pts [B,N,3] or [total,3] f32 cuda
    (+ offsets [B+1] i64 cuda).  Returns (centred points f32, centres [B,3] f64); `out` may be `pts` itself."""
    L = _native.lib()
    assert pts.is_cuda and pts.dtype == torch.float32
    pts = pts.contiguous()
    if offsets is None:
        assert pts.dim() == 3, "uniform clouds: pts must be [B,N,3]"
        B, maxn = pts.shape[0], pts.shape[1]
    else:
        assert offsets.is_cuda and offsets.dtype == torch.int64 and max_cloud_points is not None
        B, maxn = offsets.numel() - 1, int(max_cloud_points)
    out = torch.empty_like(pts) if out is None else out
    centers = torch.empty((B, 3), dtype=torch.float64, device=pts.device)
    with torch.cuda.device(pts.device):
        _native.check(L.dexgg_center_clouds(_native.ptr(pts), _native.ptr(offsets), B, maxn, pts.numel() // 3, _native.ptr(out),
                                            _native.ptr(centers), _native.stream_ptr(pts.device)))
    return out, centers


class bps_torch:
    """Drop-in for `bps_torch.bps.bps_torch` restricted to what the reference uses: a custom basis and 'dists'
    (plus 'ids'/'deltas', which upstream also returns on request)."""

    def __init__(self, bps_type="random_uniform", n_bps_points=1024, radius=1.0, n_dims=3, random_seed=13, custom_basis=None,
                 device=None, **kwargs):
        if device is None:
            device = "cuda"
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("dexgg bps_torch runs on a B200 only (no CPU fallback)")
        if custom_basis is not None:
            basis = torch.as_tensor(to_np(custom_basis), dtype=torch.float32)
        elif bps_type == "random_uniform":
            # uniform in a ball, as BPSEncoder.generate_new_bps uses it (bps_encoder.py:22-26)
            rng = np.random.default_rng(random_seed)
            x = rng.standard_normal((n_bps_points, n_dims))
            x /= np.linalg.norm(x, axis=1, keepdims=True)
            x *= rng.uniform(0, 1, (n_bps_points, 1)) ** (1.0 / n_dims) * radius
            basis = torch.as_tensor(x, dtype=torch.float32)
        else:
            raise NotImplementedError("only custom_basis / random_uniform bases are supported")
        assert basis.dim() == 2 and basis.shape[1] == 3
        self.bps = basis.reshape(1, -1, 3).to(self.device)
        self._ws = _Workspace()

    def encode(self, x, feature_type=("dists",), x_features=None, custom_basis=None, **kwargs):
        x = torch.as_tensor(to_np(x) if not torch.is_tensor(x) else x).to(dtype=torch.float32, device=self.device)
        if x.dim() == 2:
            x = x.unsqueeze(0)
        basis = self.bps[0] if custom_basis is None else torch.as_tensor(to_np(custom_basis), dtype=torch.float32,
                                                                      device=self.device).reshape(-1, 3)
        dists, idx = bps_encode_device(basis, x, ws=self._ws)
        out = {}
        if "dists" in feature_type:
            out["dists"] = dists
        if "ids" in feature_type or "closest" in feature_type or "deltas" in feature_type:
            ids = idx.long()
            out["ids"] = ids
            closest = torch.gather(x, 1, ids.clamp_min(0).unsqueeze(-1).expand(-1, -1, 3))
            if "closest" in feature_type:
                out["closest"] = closest
            if "deltas" in feature_type:
                out["deltas"] = closest - basis.unsqueeze(0)
        return out


class BPSEncoder:
    """DexGanGrasp.data.bps_encoder.BPSEncoder (bps_encoder.py:12-55): same constructor `(cfg, bps_path, vis=False)` (the
    device is cuda:cfg["gpu_ids"][0], :16-17), same static `generate_new_bps(base_path)`, same
    `encode_pcd_with_bps(pcd) -> np.float32 [4096]`, same `min < 0.1` assertion.  `bps_path=None` loads the basis shipped with
    this package (the reference's models/basis_point_set.npy); `vis` is accepted and ignored (Open3D windows are out of scope:
    no graphics stack)."""

    def __init__(self, cfg=None, bps_path=None, vis=False):
        if bps_path is None:
            bps_path = BASIS_PATH
        self.bps_np = np.load(bps_path)
        gpu = cfg["gpu_ids"][0] if (cfg is not None and "gpu_ids" in cfg and len(cfg["gpu_ids"])) else torch.cuda.current_device()
        self.device = torch.device("cuda:{}".format(gpu))
        self.bps = bps_torch(custom_basis=self.bps_np, device=self.device)
        self.VISUALIZE = False

    @staticmethod
    def generate_new_bps(base_path, n_points=4096, radius=0.2):
        """bps_encoder.py:22-26: a new uniform-in-ball basis saved as <base_path>/basis_point_set.npy."""
        b = bps_torch(bps_type="random_uniform", n_bps_points=n_points, radius=radius, n_dims=3)
        np.save(os.path.join(base_path, "basis_point_set.npy"), to_np(b.bps.squeeze()))

    def encode_pcd_with_bps(self, obj_pcd):
        pts = np.asarray(obj_pcd.points if hasattr(obj_pcd, "points") else obj_pcd)
        pc_tensor = torch.from_numpy(np.ascontiguousarray(pts))
        enc = self.bps.encode(pc_tensor)["dists"]
        enc_np = enc.cpu().detach().numpy()
        assert enc_np.min() < 0.1, "The pcd might not be in centered in origin !"
        return np.squeeze(enc_np)
