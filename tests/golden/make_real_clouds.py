"""Full-size real-object fixtures: the xyz columns of ALL seven segmented clouds of
/root/reference/data/real_objects/object/*.pcd, un-reduced (15 009 - 72 641 points, float32 exactly as stored), with the
stored encodings data/real_objects/bps/*.npy and the oracle's nearest-neighbour indices, so that config C1 runs on the GPU box
on the clouds the reference ships (the box has no /root/reference).  Run in the build container:
    python tests/golden/make_real_clouds.py
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
REF = "/root/reference"

from oracle import pcd  # noqa: E402


def main():
    lib = ctypes.CDLL(os.path.join(REPO, "oracle", "libdexgg_oracle.so"))
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    basis = np.load(os.path.join(REF, "models/basis_point_set.npy")).astype(np.float32)
    names = sorted(f[:-4] for f in os.listdir(os.path.join(REF, "data/real_objects/object")) if f.endswith(".pcd"))
    out = {"names": np.array(names)}
    for name in names:
        pts = pcd.read_pcd_xyz(os.path.join(REF, "data/real_objects/object", name + ".pcd"))
        gold = np.load(os.path.join(REF, "data/real_objects/bps", name + ".npy"))
        d = np.empty(4096, np.float32)
        idx = np.empty(4096, np.int32)
        lib.dexgg_oracle_bps_encode_mt(vp(basis), 4096, vp(pts), len(pts), vp(d), vp(idx), 8)
        assert np.array_equal(d, gold[0]), name  # the oracle reproduces the reference's stored encoding bit for bit
        out[name + "/cloud"] = pts
        out[name + "/bps"] = gold
        out[name + "/idx"] = idx
        print(f"{name}: {len(pts)} points, oracle == stored encoding")
    np.savez_compressed(os.path.join(HERE, "real_clouds_full.npz"), **out)


if __name__ == "__main__":
    main()
