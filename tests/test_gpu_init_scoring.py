"""Row N2: Initializer::CheckHomography / CheckFundamental on the device against the oracle
(Initialization/Initializer.cpp:268-438, pinned to the reference's own function bodies by
tests/test_init_scoring_oracle.py).  Bit-exact: the float32 scores (an ordered running sum; NaN scores of degenerate hypotheses compare as NaN), the inlier flags of every
hypothesis, the list length and the RANSAC selection (:205-209, :257-261), including degenerate hypotheses
(zero matrix, overflow, NaN, exact ties)."""
import numpy as np
import pytest

from init_util import golden_scoring, hypotheses, same_scores, scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    from orb_slam_tracking_b200 import frontend
    return frontend


class _F:  # a frame as the Initializer sees it: undistorted keys only
    def __init__(self, ex, k):
        self.ex, self.mvKeysUn = ex, k


@pytest.mark.parametrize("seed,sigma,nh", [(1, 1.0, 200), (2, 1.0, 64), (3, 0.7, 300), (4, 2.5, 1), (5, 1.0, 8)])
def test_check_models_host_pair(fe, oracle, seed, sigma, nh):
    E = fe.ORBextractor(1000, 1.2, 8, 20, 7, 640, 480)
    k1, k2, m12, H = scene(seed)
    H21, H12, F21 = hypotheses(seed, H, nh)
    I = fe.Initializer(_F(E, k1), sigma, nh)
    with np.errstate(all="ignore"):
        sO, iO = oracle.check_homography(k1, k2, m12, H21, H12, sigma)
        sG, iG, bG = I.CheckHomography(_F(E, k2), m12, H21, H12)
        assert same_scores(sG, sO) and np.array_equal(iG, iO) and bG == oracle.ransac_best(sO)
        sO, iO = oracle.check_fundamental(k1, k2, m12, F21, sigma)
        sG, iG, bG = I.CheckFundamental(_F(E, k2), m12, F21)
        assert same_scores(sG, sO) and np.array_equal(iG, iO) and bG == oracle.ransac_best(sO)
    if nh >= 8:
        assert bG != 7   # hypothesis 7 repeats hypothesis 0: the earlier one wins a tie
    E.close()


def test_check_models_edge_cases(fe, oracle):
    E = fe.ORBextractor(1000, 1.2, 8, 20, 7, 640, 480)
    k1, k2, m12, H = scene(9)
    H21, H12, F21 = hypotheses(9, H, 4)
    none = np.full(len(k1), -1, np.int32)
    I = fe.Initializer(_F(E, k1))
    s, i, b = I.CheckHomography(_F(E, k2), none, H21, H12)
    assert s.tolist() == [0, 0, 0, 0] and i.shape == (4, 0) and b == -1
    s, i, b = I.CheckFundamental(_F(E, k2), none, F21)
    assert s.tolist() == [0, 0, 0, 0] and b == -1
    bad = m12.copy()
    bad[np.argmax(m12 >= 0)] = len(k2)   # an index past the second frame is an argument error, not a wild read
    with pytest.raises(Exception):
        I.CheckHomography(_F(E, k2), bad, H21, H12)
    E.close()


def test_stream_score_models(fe, oracle):
    """The resident variant: frames and vnMatches12 stay on the device between the matcher and the scoring."""
    from orb_slam_tracking_b200 import synth
    W, H, nf, B, nh = 640, 480, 1000, 8, 96
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H, max_batch=B)
    E.set_camera(K4, D4)
    S = fe.Stream(E, 0.9, True, 100)
    S.upload(synth.frames(B, W, H, first_pair=40))
    S.run(B)
    out = S.download(B)
    nP = B // 2
    H21 = np.zeros((nP, nh, 9), np.float32); H12 = np.zeros_like(H21); F21 = np.zeros_like(H21)
    rng = np.random.default_rng(3)
    for p in range(nP):
        # hypotheses around the translation that the matches suggest
        m = out["matches12"][p][: out["n"][2 * p]]
        a, b = out["kps_un"][2 * p][: len(m)][m >= 0], out["kps_un"][2 * p + 1][m[m >= 0]]
        t = np.array([np.median(b["x"] - a["x"]), np.median(b["y"] - a["y"])]) if len(a) else np.zeros(2)
        Ht = np.array([[1, 0, t[0]], [0, 1, t[1]], [0, 0, 1.0]])
        H21[p], H12[p], F21[p] = hypotheses(100 + p, Ht, nh)
    res = S.score_models(nP, H21, H12, F21, sigma=1.0)
    with np.errstate(all="ignore"):
        for p in range(nP):
            n1 = out["n"][2 * p]
            k1, k2, m12 = out["kps_un"][2 * p][:n1], out["kps_un"][2 * p + 1][: out["n"][2 * p + 1]], out["matches12"][p][:n1]
            for tag, (sO, iO) in (("h", oracle.check_homography(k1, k2, m12, H21[p], H12[p])),
                                  ("f", oracle.check_fundamental(k1, k2, m12, F21[p]))):
                N = iO.shape[1]
                assert res["n_list"][p] == N
                assert same_scores(res["scores_" + tag][p], sO), (p, tag)
                b = oracle.ransac_best(sO)
                assert res["best_" + tag][p] == b
                want = np.zeros(E.cap, np.uint8)
                if b >= 0:
                    want[:N] = iO[b]
                assert np.array_equal(res["inliers_" + tag][p], want), (p, tag)
    assert (res["n_list"] > 20).all() and (res["best_h"] >= 0).all()
    # one model at a time
    only = S.score_models(nP, F21=F21)
    assert only["scores_h"] is None and same_scores(only["scores_f"], res["scores_f"])
    E.close()


@pytest.mark.parametrize("tag,sigma", [("s10", 1.0), ("s5", 0.5)])
def test_golden_demo_pair(fe, tag, sigma):
    """CUDA against the committed outputs of the reference's own scoring bodies on its demo pair (no oracle involved)."""
    g = golden_scoring()
    E = fe.ORBextractor(2000, 1.2, 8, 0, 0, 752, 480)
    I = fe.Initializer(_F(E, g["kps_un0"]), sigma, len(g["H21"]))
    s, i, b = I.CheckHomography(_F(E, g["kps_un1"]), g["matches12"], g["H21"], g["H12"])
    assert same_scores(s, g["scoresH_" + tag]) and np.array_equal(np.packbits(i, axis=1), g["inliersH_" + tag])
    assert b == 0 or s[b] > s[0]
    s, i, b = I.CheckFundamental(_F(E, g["kps_un1"]), g["matches12"], g["F21"])
    assert same_scores(s, g["scoresF_" + tag]) and np.array_equal(np.packbits(i, axis=1), g["inliersF_" + tag])
    E.close()
