"""Row N2: the oracle's CheckHomography / CheckFundamental against the reference's own function
bodies (oracle/_ref, cut from Initialization/Initializer.cpp:268-438 at build time) -- bit-exact scores and inliers."""
import numpy as np
import pytest

from oracle import orb_oracle as orc
from oracle import ref_orb
from init_util import golden_scoring, hypotheses, same_scores, scene

needs_ref = pytest.mark.skipif(not (ref_orb.available() and hasattr(ref_orb.lib(), "ref_check_homography")),
                               reason="oracle/_ref not built (no /root/reference on this box)")


@needs_ref
@pytest.mark.parametrize("seed,sigma", [(1, 1.0), (2, 1.0), (3, 0.7), (4, 2.5)])
def test_oracle_scoring_matches_reference(seed, sigma):
    k1, k2, m12, H = scene(seed)
    H21, H12, F21 = hypotheses(seed, H, 64)
    with np.errstate(all="ignore"):
        sO, iO = orc.check_homography(k1, k2, m12, H21, H12, sigma)
        sR, iR = ref_orb.check_homography(k1, k2, m12, H21, H12, sigma)
        assert sO.tobytes() == sR.tobytes() and np.array_equal(iO, iR)
        assert iO.shape[1] == int((m12 >= 0).sum()) and iO[0].sum() > 50   # the near-true model explains the inlier matches
        sO, iO = orc.check_fundamental(k1, k2, m12, F21, sigma)
        sR, iR = ref_orb.check_fundamental(k1, k2, m12, F21, sigma)
        assert sO.tobytes() == sR.tobytes() and np.array_equal(iO, iR)


def test_ransac_best_rule():
    assert orc.ransac_best(np.array([0, 0, 0], np.float32)) == -1
    assert orc.ransac_best(np.array([1, 5, 5, 2], np.float32)) == 1
    assert orc.ransac_best(np.array([np.nan, 3, np.nan, 3], np.float32)) == 1
    assert orc.ransac_best(np.array([-1, -2], np.float32)) == -1


def test_no_matches():
    k1, k2, m12, H = scene(5)
    m12[:] = -1
    H21, H12, F21 = hypotheses(5, H, 4)
    s, i = orc.check_homography(k1, k2, m12, H21, H12)
    assert s.tolist() == [0, 0, 0, 0] and i.shape == (4, 0)
    s, i = orc.check_fundamental(k1, k2, m12, F21)
    assert s.tolist() == [0, 0, 0, 0]


@pytest.mark.parametrize("tag,sigma", [("s10", 1.0), ("s5", 0.5)])
def test_oracle_matches_golden_demo_pair(tag, sigma):
    """tests/golden/init_scoring_demo.npz: the reference's own scoring bodies on its demo pair (tools/gen_golden_init.py)."""
    g = golden_scoring()
    with np.errstate(all="ignore"):
        s, i = orc.check_homography(g["kps_un0"], g["kps_un1"], g["matches12"], g["H21"], g["H12"], sigma)
        assert same_scores(s, g["scoresH_" + tag]) and np.array_equal(np.packbits(i, axis=1), g["inliersH_" + tag])
        s, i = orc.check_fundamental(g["kps_un0"], g["kps_un1"], g["matches12"], g["F21"], sigma)
        assert same_scores(s, g["scoresF_" + tag]) and np.array_equal(np.packbits(i, axis=1), g["inliersF_" + tag])
