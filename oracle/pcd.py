"""TEST INFRASTRUCTURE (oracle/): minimal binary PCD v0.7 reader.

The reference reads its real-object clouds with Open3D
(`dexgangrasp_offline.py:134-147`, `DexGanGrasp/data/bps_encoder.py:28-31`:
`np.asarray(obj_pcd.points)`, float64 [N,3]).  Open3D is absent here, so this
reads the same files (`FIELDS x y z [normal_x normal_y normal_z] rgb`, all
4-byte, `DATA binary`) and returns xyz exactly as stored (float32), which is
what bps_torch's `to_tensor` cast recovers from Open3D's float64 copy.
"""
import numpy as np


def read_pcd_xyz(path):
    with open(path, "rb") as f:
        raw = f.read()
    header_end = raw.index(b"DATA binary\n") + len(b"DATA binary\n")
    header = raw[:header_end].decode("ascii").splitlines()
    meta = {}
    for line in header:
        if line.startswith("#") or not line.strip():
            continue
        k, *v = line.split()
        meta[k] = v
    fields = meta["FIELDS"]
    sizes = [int(s) for s in meta["SIZE"]]
    counts = [int(c) for c in meta["COUNT"]]
    assert all(s == 4 for s in sizes) and all(c == 1 for c in counts), "only 4-byte scalar fields"
    assert fields[:3] == ["x", "y", "z"] and meta["TYPE"][:3] == ["F", "F", "F"]
    n = int(meta["POINTS"][0])
    rec = np.frombuffer(raw, dtype=np.float32, count=n * len(fields), offset=header_end)
    return np.ascontiguousarray(rec.reshape(n, len(fields))[:, :3])


class PointCloudLike:
    """Stand-in for an Open3D PointCloud: exposes `.points` (float64 [N,3])."""

    def __init__(self, xyz):
        self.points = np.asarray(xyz, dtype=np.float64)
