"""The brute-force matcher (csrc/match_bf.cu): SearchForInitialization with the Hamming matrix sharded by query rows,
resolved in the reference's order from per-query top lists.  Bit-exact against the oracle (Features/ORBmatcher.cpp:11-183)
on every path of launch_match, at 3840x2160 / 8000 keypoints (BASELINE.json configs[4]), with the rows split as the
multi-GPU job splits them, and on adversarial frames where many queries compete for the same candidates."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    from orb_slam_tracking_b200 import frontend
    return frontend


@pytest.fixture(scope="module")
def synth():
    from orb_slam_tracking_b200 import synth
    return synth


def _frames(fe, oracle, synth, W, H, nf, ini, mn, pair_index):
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, ini, mn, W, H)
    bounds = E.set_camera(K4, D4)
    R = oracle.Extractor(nf, 1.2, 8, ini, mn)
    imgs = synth.frame_pair(pair_index, W, H)
    G = [fe.Frame(E, im) for im in imgs]
    O = []
    for im in imgs:
        k, d, _ = R(im)
        O.append(oracle.Frame(k, d, K4, D4, bounds))
    return E, G, O


# the 3200-feature case: more grid entries than k_match_cta's scratch view of them holds (its set-up walks global memory instead)
@pytest.mark.parametrize("cfg", [(640, 480, 1000, 20, 7), (752, 480, 2000, 0, 0), (1241, 376, 2000, 20, 7), (1241, 376, 3200, 0, 0)])
def test_every_matcher_path_agrees_with_the_oracle(fe, oracle, synth, cfg):
    W, H, nf, ini, mn = cfg
    E, G, O = _frames(fe, oracle, synth, W, H, nf, ini, mn, 33)
    for nnratio, ori, window in ((0.9, True, 100), (0.6, True, 100), (0.95, False, 300), (0.9, True, 8192), (0.9, True, 7)):
        rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], nnratio, ori, window)
        for path in (1, 2, 3, 0):
            E.set_option("match_path", path)
            n, m12, cnt = fe.ORBmatcher(nnratio, ori).SearchForInitialization(G[0], G[1], window)
            assert np.array_equal(m12, rm12), (path, nnratio, ori, window, int((m12 != rm12).sum()))
            assert n == rn and np.array_equal(cnt, rcnt), (path, n, rn, cnt, rcnt)
    # identical frames: every level-0 key finds itself at distance 0
    E.set_option("match_path", 3)
    n, m12, _ = fe.ORBmatcher(0.9, True).SearchForInitialization(G[0], G[0], 8192)
    rn, rm12, _ = oracle.search_for_initialization(O[0], O[0], 0.9, True, 8192)
    assert n == rn and np.array_equal(m12, rm12)
    E.close()


def test_uhd_8000_keypoints(fe, oracle, synth):
    """BASELINE.json configs[4]: 3840x2160, 8000 keypoints, window 100 and the all-pairs window."""
    W, H, nf = 3840, 2160, 8000
    E, G, O = _frames(fe, oracle, synth, W, H, nf, 20, 7, 2)
    assert G[0].N == O[0].N == 8000
    for window in (100, 8192):
        rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], 0.9, True, window)
        for path in (0, 3, 1):
            E.set_option("match_path", path)
            n, m12, cnt = fe.ORBmatcher(0.9, True).SearchForInitialization(G[0], G[1], window)
            assert np.array_equal(m12, rm12), (window, path, int((m12 != rm12).sum()))
            assert n == rn and np.array_equal(cnt, rcnt), (window, path)
    E.close()


def test_rows_sharded_like_the_multi_gpu_job(fe, oracle, synth):
    """orbx_match_rows on row blocks (what each rank computes) + orbx_match_resolve == SearchForInitialization."""
    from orb_slam_tracking_b200 import shard
    E, G, O = _frames(fe, oracle, synth, 752, 480, 2000, 0, 0, 34)
    M = fe.ORBmatcher(0.9, True)
    for window in (8192, 100):
        rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], 0.9, True, window)
        _, nq = M.match_rows(G[0], G[1], window, 0, 0)
        assert nq == int((O[0].keys_un()["octave"] <= 0).sum())
        for world in (1, 3, 8):
            blocks = [M.match_rows(G[0], G[1], window, *shard.shard_rows(nq, world, r))[0] for r in range(world)]
            rows = np.concatenate(blocks)
            assert len(rows) == nq
            n, m12, cnt = M.match_resolve(G[0], G[1], rows, window)
            assert n == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt), (window, world)
    E.close()


def test_contended_candidates(fe, oracle):
    """Few distinct descriptors on many keypoints: distance ties everywhere, and most queries find their best candidates
    already taken by earlier queries (the lists run out and the ordered rescan takes over)."""
    W, H, nf = 640, 480, 3000
    K4, D4 = np.array([500, 500, 320, 240], np.float32), np.zeros(4, np.float32)
    E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H)
    bounds = E.set_camera(K4, D4)
    rng = np.random.default_rng(11)
    from orb_slam_tracking_b200.capi import KP_DTYPE
    codes = rng.integers(0, 256, (12, 32), dtype=np.uint8)

    def frame(n, seed):
        r = np.random.default_rng(seed)
        k = np.zeros(n, KP_DTYPE)
        k["x"] = np.floor(r.uniform(20, W - 20, n)); k["y"] = np.floor(r.uniform(20, H - 20, n))
        k["size"] = 31; k["angle"] = r.uniform(0, 360, n).astype(np.float32); k["response"] = 30
        k["octave"] = (r.random(n) < 0.2).astype(np.int32); k["class_id"] = -1
        d = codes[r.integers(0, len(codes), n)].copy()
        flip = r.random(n) < 0.5                       # half of them a few bits away from their code
        d[flip, 0] ^= r.integers(0, 256, int(flip.sum()), dtype=np.uint8)
        return k, d

    for n1, n2 in ((2500, 2500), (3000, 40), (40, 3000)):
        (k1, d1), (k2, d2) = frame(n1, 1), frame(n2, 2)
        G = [fe.Frame(E, kps=k1, desc=d1), fe.Frame(E, kps=k2, desc=d2)]
        O = [oracle.Frame(k1, d1, K4, D4, bounds), oracle.Frame(k2, d2, K4, D4, bounds)]
        for window in (8192, 200, 60):
            for nnratio in (0.9, 1.0):
                rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], nnratio, True, window)
                for path in (3, 1):
                    E.set_option("match_path", path)
                    n, m12, cnt = fe.ORBmatcher(nnratio, True).SearchForInitialization(G[0], G[1], window)
                    assert np.array_equal(m12, rm12), (n1, n2, window, nnratio, path, int((m12 != rm12).sum()))
                    assert n == rn and np.array_equal(cnt, rcnt), (n1, n2, window, nnratio, path, n, rn, cnt, rcnt)
    E.close()


def test_raw_all_pairs_hamming(fe):
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (1900, 32), dtype=np.uint8)
    b[100:140] = a[:40]            # exact matches
    b[300:340] = a[:40]            # ... twice: the first one in index order wins, the second distance is 0 too
    b[500] = b[501]                # tied candidates
    E = fe.ORBextractor(100, 1.2, 2, 20, 7, 640, 480)
    bi, bd, sd, ms = E.hamming_best2(a, b, repeats=2)
    D = np.unpackbits(a[:, None, :] ^ b[None, :, :], axis=2).sum(2)
    assert np.array_equal(bi, D.argmin(1))
    assert np.array_equal(bd, D.min(1))
    assert np.array_equal(sd, np.sort(D, 1)[:, 1])
    assert ms > 0
    bi, bd, sd, _ = E.hamming_best2(a[:5], b[:1])
    assert (bi == 0).all() and (sd == -1).all()
    E.close()
