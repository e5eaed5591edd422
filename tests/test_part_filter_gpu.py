"""Part-conditioned ranking on the device (csrc/part_rank.cu through the C ABI) against the oracle."""
import numpy as np
import pytest

from oracle import part_filter_oracle as po
from test_oracle_part_filter import _case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,n,m", [(0, 200, 3000), (1, 1, 10), (2, 4097, 100000), (3, 16384, 50)])
def test_ranking_matches_oracle(seed, n, m):
    from dex_gan_grasp_b200 import part_filter
    grasps, part, center = _case(seed, n, m)
    idx, mean = part_filter.filter_grasps_given_part({k: v.astype(np.float32) for k, v in grasps.items()}, part.astype(np.float32), center)
    ridx, rmean, rmins = po.filter_grasps_given_part(grasps, part, center)
    assert np.allclose(mean, rmean, atol=2e-7)   # f32 inputs, f64 accumulation
    assert sorted(idx.tolist()) == list(range(n))  # a permutation
    # same order wherever the oracle's distances are separated by more than the f32 resolution of the score
    assert np.allclose(rmins[idx], rmins[ridx], rtol=1e-5, atol=1e-9)
    sep = np.diff(rmins[ridx]) > 1e-5 * rmins[ridx][1:] + 1e-9
    firm = np.concatenate([[True], sep]) & np.concatenate([sep, [True]])
    assert np.array_equal(idx[firm], ridx[firm])
    a = part_filter.sort_grasps({k: v.copy() for k, v in grasps.items()}, idx, 30)
    assert a["transl"].shape[0] == min(30, n)
