"""TEST INFRASTRUCTURE (imported by tests/ only).  CPU restatement of the centring the reference's callers do with Open3D before the
BPS encoder: `pc_center = obj_pcd.get_center(); obj_pcd.translate(-pc_center)` (dexgangrasp_real_robot.py:133-134,
DexGanGrasp/data/bps_encoder.py:69).  Open3D is a C++ dependency that is not in this image (out of scope), so this is
its published semantics restated - PointCloud points are float64, get_center() is their mean, translate() adds in float64, and
bps_torch turns the result into float32 - and it is NOT pinned by running Open3D here ("parity unpinned" for this one helper).
Sanity only: the shipped .pcd fixtures were stored after that centring (and a later outlier removal), and their float64 mean is
~1e-3 m (tests/test_center_clouds.py)."""
import numpy as np


def center_cloud(points_f32):
    """points [N,3] float32 -> (centred float32 [N,3], centre float64 [3])"""
    p = np.asarray(points_f32, dtype=np.float32).astype(np.float64)
    c = p.mean(axis=0) if len(p) else np.zeros(3)
    return (p - c).astype(np.float32), c
