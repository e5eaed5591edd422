"""Shim of the third-party `bps_torch` package (git+https://github.com/otaheri/bps_torch, unpinned in the reference's
README.md:41) restricted to what the reference calls: `bps_torch.bps.bps_torch(custom_basis=..., device=...).encode(x)['dists']`
(DexGanGrasp/data/bps_encoder.py:19,33; dexgangrasp_real_robot.py:21,101,139; dexafford_prompt.py:19,123,177) and
`bps_torch.utils.to_np` (bps_encoder.py:5).  The nearest-neighbour search runs in the sm_100a kernel of libdexgg; the
`chamfer_distance` extension upstream needs is not used."""
