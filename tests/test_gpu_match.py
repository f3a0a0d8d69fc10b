"""Frame glue (undistort + 64x48 grid), GetFeaturesInArea, SearchForInitialization and the
device-resident stream path against the oracle (reference: SlamTypes/Frame.cpp:40-206,
Features/ORBmatcher.cpp:11-183).  Bit-exact: undistorted coordinates, grid lists, candidate order,
vnMatches12, the returned count (with its double-decrement quirk) and the three printed counters."""
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


def _pair(fe, oracle, synth, W, H, nf, ini, mn, pair_index, K4=None, D4=None):
    if K4 is None:
        K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, ini, mn, W, H)
    bounds = E.set_camera(K4, D4)
    assert np.array_equal(bounds, oracle.image_bounds(W, H, K4, D4))
    R = oracle.Extractor(nf, 1.2, 8, ini, mn)
    imgs = synth.frame_pair(pair_index, W, H)
    G, O = [], []
    for im in imgs:
        G.append(fe.Frame(E, im))
        k, d, _ = R(im)
        O.append(oracle.Frame(k, d, K4, D4, bounds))
    return E, G, O


@pytest.mark.parametrize("shape", [(752, 480, 1200), (640, 480, 1000), (1241, 376, 2000)])
def test_frame_glue(fe, oracle, synth, shape):
    W, H, nf = shape
    E, G, O = _pair(fe, oracle, synth, W, H, nf, 20, 7, 21)
    for g, o in zip(G, O):
        ku = o.keys_un()
        assert g.N == o.N
        for f in ku.dtype.names:
            assert np.array_equal(g.mvKeysUn[f].view(np.uint32), ku[f].view(np.uint32)), f
        off, idx = o.grid()
        assert np.array_equal(g.grid_offsets, off)
        assert np.array_equal(g.grid_indices, idx)
    E.close()


def test_features_in_area(fe, oracle, synth):
    E, G, O = _pair(fe, oracle, synth, 752, 480, 1200, 20, 7, 22)
    rng = np.random.default_rng(1)
    g, o = G[1], O[1]
    qs = [(float(x), float(y), float(r), lv) for x, y, r, lv in
          zip(rng.uniform(-100, 900, 60), rng.uniform(-100, 600, 60), rng.uniform(5, 300, 60), rng.integers(-1, 4, 60))]
    qs += [(376.0, 240.0, 8192.0, 0), (0.0, 0.0, 100.0, -1), (5000.0, 5000.0, 10.0, 0), (-5000.0, 0.0, 10.0, 0)]
    for x, y, r, lv in qs:
        a = g.GetFeaturesInArea(x, y, r, int(lv), int(lv))
        b = o.features_in_area(x, y, r, int(lv), int(lv))
        assert np.array_equal(a, b), (x, y, r, lv)
    E.close()


@pytest.mark.parametrize("cfg", [(640, 480, 1000, 20, 7), (752, 480, 2000, 0, 0), (752, 480, 1200, 20, 7),
                                 (1241, 376, 2000, 20, 7)])
def test_search_for_initialization(fe, oracle, synth, cfg):
    W, H, nf, ini, mn = cfg
    for pair_index in (30, 31):
        E, G, O = _pair(fe, oracle, synth, W, H, nf, ini, mn, pair_index)
        for nnratio, ori, window in ((0.9, True, 100), (0.6, True, 100), (0.9, False, 50), (0.9, True, 8192)):
            n, m12, cnt = fe.ORBmatcher(nnratio, ori).SearchForInitialization(G[0], G[1], window)
            rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], nnratio, ori, window)
            assert np.array_equal(m12, rm12), (nnratio, ori, window, int((m12 != rm12).sum()))
            assert n == rn and np.array_equal(cnt, rcnt), (n, rn, cnt, rcnt)
        # identical frames: every level-0 key finds itself at distance 0
        n, m12, _ = fe.ORBmatcher(0.9, True).SearchForInitialization(G[0], G[0], 100)
        rn, rm12, _ = oracle.search_for_initialization(O[0], O[0], 0.9, True, 100)
        assert n == rn and np.array_equal(m12, rm12)
        E.close()


def test_empty_frames(fe, oracle):
    W, H = 640, 480
    E = fe.ORBextractor(1000, 1.2, 8, 20, 7, W, H)
    E.set_camera(np.array([500, 500, 320, 240], np.float32), np.zeros(4, np.float32))
    flat = np.full((H, W), 50, np.uint8)
    F = fe.Frame(E, flat)
    assert F.N == 0 and F.grid_offsets[-1] == 0
    n, m12, cnt = fe.ORBmatcher(0.9, True).SearchForInitialization(F, F, 100)
    assert n == 0 and len(m12) == 0 and not cnt.any()
    E.close()


def test_descriptor_distance(fe, oracle):
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    b[:50] = a[:50]
    ref = np.array([oracle.descriptor_distance(x, y) for x, y in zip(a, b)])
    host = np.array([fe.ORBmatcher.DescriptorDistance(x, y) for x, y in zip(a, b)])
    assert np.array_equal(ref, host)
    from orb_slam_tracking_b200.capi import lib, ptr, check
    E = fe.ORBextractor(100, 1.2, 2, 20, 7, 640, 480)
    out = np.zeros(500, np.int32)
    check(lib().orbx_descriptor_distance_batch(E._h, ptr(a), ptr(b), 500, ptr(out)), E._h)
    assert np.array_equal(out, ref)
    E.close()


@pytest.mark.parametrize("cfg", [(640, 480, 1000, 20, 7, 8), (752, 480, 1200, 20, 7, 6)])
def test_stream_path(fe, oracle, synth, cfg):
    W, H, nf, ini, mn, B = cfg
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, ini, mn, W, H, max_batch=B)
    bounds = E.set_camera(K4, D4)
    R = oracle.Extractor(nf, 1.2, 8, ini, mn)
    imgs = synth.frames(B, W, H, first_pair=40)
    S = fe.Stream(E, 0.9, True, 100)
    S.upload(imgs)
    S.run(B)
    out = S.download(B)
    frames = []
    for i in range(B):
        k, d, _ = R(imgs[i])
        n = out["n"][i]
        assert n == len(k)
        assert np.array_equal(out["kps"][i, :n].view(np.uint8), k.view(np.uint8)), i
        assert np.array_equal(out["desc"][i, :n], d), i
        F = oracle.Frame(k, d, K4, D4, bounds)
        assert np.array_equal(out["kps_un"][i, :n].view(np.uint8), F.keys_un().view(np.uint8)), i
        frames.append(F)
    for p in range(B // 2):
        rn, rm12, rcnt = oracle.search_for_initialization(frames[2 * p], frames[2 * p + 1], 0.9, True, 100)
        assert out["nmatches"][p] == rn, p
        assert np.array_equal(out["matches12"][p, :len(rm12)], rm12), p
        assert np.array_equal(out["counters"][p], rcnt), p
    # size-independent property at full batch: re-running the same batch is idempotent
    S.run(B)
    out2 = S.download(B)
    for k in out:
        assert np.array_equal(out[k].view(np.uint8), out2[k].view(np.uint8)), k
    E.close()


def test_matcher_chunking_and_wide_windows(fe, oracle, synth):
    """Windows that cover the image (brute-force mode of config 5) make every level-0 key a candidate of
    every query, so the entry buffer of k_match_cta is filled in several chunks."""
    W, H, nf = 752, 480, 2000
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, 0, 0, W, H)   # thresholds 0/0: full quotas, tie-heavy distances
    bounds = E.set_camera(K4, D4)
    R = oracle.Extractor(nf, 1.2, 8, 0, 0)
    imgs = synth.frames(2, W, H, first_pair=21)
    G = [fe.Frame(E, im) for im in imgs]
    O = []
    for im in imgs:
        k, d, _ = R(im)
        O.append(oracle.Frame(k, d, K4, D4, bounds))
    for w in (8192, 1000, 300, 100, 7, 1):
        for ratio, ori in ((0.9, True), (0.6, True), (0.95, False)):
            n, m12, cnt = fe.ORBmatcher(ratio, ori).SearchForInitialization(G[0], G[1], w)
            rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], ratio, ori, w)
            assert n == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt), (w, ratio, ori)
    E.close()


def test_serial_matcher_fallback():
    """k_match (one warp per pair) is the fallback for keypoint capacities beyond the shared-memory tables of
    k_match_cta; ORBX_MATCH_SERIAL forces it so that it stays covered.  Own process: the hook is read once."""
    import os
    import subprocess
    import sys
    code = (
        "import numpy as np\n"
        "from oracle import orb_oracle as O\n"
        "from orb_slam_tracking_b200 import frontend as fe, synth\n"
        "W, H, nf = 640, 480, 1000\n"
        "K4, D4 = synth.camera_for(W, H)\n"
        "E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H)\n"
        "b = E.set_camera(K4, D4)\n"
        "R = O.Extractor(nf, 1.2, 8, 20, 7)\n"
        "imgs = synth.frames(2, W, H, first_pair=9)\n"
        "G = [fe.Frame(E, im) for im in imgs]\n"
        "F = [O.Frame(*R(im)[:2], K4, D4, b) for im in imgs]\n"
        "for w in (100, 640):\n"
        "    n, m, c = fe.ORBmatcher(0.9, True).SearchForInitialization(G[0], G[1], w)\n"
        "    rn, rm, rc = O.search_for_initialization(F[0], F[1], 0.9, True, w)\n"
        "    assert n == rn and np.array_equal(m, rm) and np.array_equal(c, rc), w\n"
        "print('serial ok')\n")
    env = dict(os.environ, ORBX_MATCH_SERIAL="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True)
    assert out.returncode == 0 and "serial ok" in out.stdout, out.stdout + out.stderr


@pytest.mark.parametrize("chunk", [2, 4, 6, 0])
def test_stream_process_chunked(fe, oracle, synth, chunk):
    """orbx_stream_process (chunked, three streams) returns exactly what upload/run/download returns."""
    W, H, nf, B = 640, 480, 1000, 10
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H, max_batch=B)
    E.set_camera(K4, D4)
    imgs = synth.frames(B, W, H, first_pair=60)
    S = fe.Stream(E, 0.9, True, 100)
    S.upload(imgs)
    S.run(B)
    ref = S.download(B)
    for n in (B, B - 1, 1):
        out = S.process(imgs[:n], chunk=chunk)
        for k in ("n", "kps", "kps_un", "desc"):
            assert np.array_equal(out[k][:n].view(np.uint8), ref[k][:n].view(np.uint8)), (k, n)
        for k in ("nmatches", "counters", "matches12"):
            assert np.array_equal(out[k][:n // 2], ref[k][:n // 2]), (k, n)
    E.close()


def test_stream_process_async_two_in_flight(fe, oracle, synth):
    """orbx_stream_process_async: two steps in flight (upload of one over the kernels / download of the other) return
    exactly what the synchronous call returns, also when the schedule changes between steps and after other calls."""
    W, H, nf, B = 640, 480, 1000, 12
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H, max_batch=B)
    E.set_camera(K4, D4)
    S = fe.Stream(E, 0.9, True, 100)
    sets = [synth.frames(B, W, H, first_pair=70 + 6 * i) for i in range(4)]
    want = [S.process(x, chunk=4) for x in sets]

    def same(a, b, n):
        for k in ("n", "kps", "kps_un", "desc"):
            assert np.array_equal(a[k][:n].view(np.uint8), b[k][:n].view(np.uint8)), k
        for k in ("nmatches", "counters", "matches12"):
            assert np.array_equal(a[k][:n // 2], b[k][:n // 2]), k

    # a stream of steps, always two in flight
    order = [0, 1, 2, 3, 0, 2, 1, 3]
    tickets = []
    for i, j in enumerate(order):
        tickets.append((S.process_async(sets[j], chunk=4), j))
        if len(tickets) == 2:
            t, jj = tickets.pop(0)
            same(S.wait(t), want[jj], B)
    t, jj = tickets.pop(0)
    same(S.wait(t), want[jj], B)
    # the default schedule of the streaming form (three equal chunks), odd batch sizes included
    for n in (B, 7, 1):
        ta = S.process_async(sets[0][:n])
        tb = S.process_async(sets[1][:n])
        same(S.wait(ta), want[0], n)
        same(S.wait(tb), want[1], n)
    # a third step without waiting is refused
    t0 = S.process_async(sets[0], chunk=4)
    t1 = S.process_async(sets[1], chunk=4)
    with pytest.raises(Exception):
        S.process_async(sets[2], chunk=4)
    same(S.wait(t0), want[0], B)
    same(S.wait(t1), want[1], B)
    # a different schedule drains the pipeline first
    t0 = S.process_async(sets[0], chunk=4)
    o0 = S._flight[t0][1]
    t1 = S.process_async(sets[1][:6], chunk=2)
    same(o0, want[0], B)            # already complete: the second call had to wait for it
    same(S.wait(t1), want[1], 6)
    same(S.wait(t0), want[0], B)
    # other entry points after in-flight steps are ordered behind them
    t0 = S.process_async(sets[3], chunk=4)
    S.upload(sets[2]); S.run(B)
    got = S.download(B)
    same(got, want[2], B)
    same(S.wait(t0), want[3], B)
    S.sync()
    E.close()


def test_stream_features_in_area_batched(fe, oracle, synth):
    """radius queries with level ranges on the frames the stream path left on the device, 1000 random
    windows per frame in one launch, against Frame::GetFeaturesInArea of the oracle (order included)."""
    W, H, nf, B = 752, 480, 1200, 4
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(nf, 1.2, 8, 20, 7, W, H, max_batch=B)
    bounds = E.set_camera(K4, D4)
    R = oracle.Extractor(nf, 1.2, 8, 20, 7)
    imgs = synth.frames(B, W, H, first_pair=50)
    S = fe.Stream(E, 0.9, True, 100)
    S.upload(imgs)
    S.run(B)
    S.sync()
    O = []
    for im in imgs:
        k, d, _ = R(im)
        O.append(oracle.Frame(k, d, K4, D4, bounds))
    rng = np.random.default_rng(8)
    n = 1000
    q = np.zeros((B * n, 5), np.float32)
    q[:, 0] = rng.uniform(-150, W + 150, B * n)
    q[:, 1] = rng.uniform(-150, H + 150, B * n)
    q[:, 2] = rng.choice([3.0, 10.0, 40.0, 100.0, 350.0, 8192.0], B * n)
    lo = rng.integers(-1, 5, B * n)
    q[:, 3] = lo
    q[:, 4] = np.where(rng.random(B * n) < 0.3, -1, lo + rng.integers(0, 4, B * n))
    q[:8] = [[0, 0, 100, -1, -1], [5000, 5000, 10, 0, 0], [-5000, 0, 10, 0, 0], [376, 240, 8192, 0, 7],
             [376, 240, 8192, -1, -1], [376, 240, 1e-3, -1, -1], [376, -4000, 10, 0, 3], [376, 240, 50, 3, 1]]
    slots = np.repeat(np.arange(B, dtype=np.int32), n)
    got, cnt = S.features_in_area(slots, q)
    for i in range(B * n):
        want = O[slots[i]].features_in_area(float(q[i, 0]), float(q[i, 1]), float(q[i, 2]), int(q[i, 3]), int(q[i, 4]))
        assert cnt[i] == len(want) and np.array_equal(got[i], want), (i, q[i])
    # a small cap truncates the lists but keeps the counts
    got2, cnt2 = S.features_in_area(slots[:50], q[:50], cap=5)
    assert np.array_equal(cnt2, cnt[:50]) and all(np.array_equal(a, b[:5]) for a, b in zip(got2, got[:50]))
    E.close()


def test_frame_glue_negative_radial_factor(fe, oracle, synth):
    """A camera whose distortion model turns over inside the image (the radial factor goes negative towards the corners): OpenCV's
    undistortPoints gives up on those points and keeps the start value; bounds, undistorted keys and grid must follow the oracle
    (pinned against the OpenCV wheel in tests/test_oracle_vs_cv2.py)."""
    W, H = 640, 480
    K4, D4 = np.array([250, 250, 320, 240], np.float32), np.array([-0.6, 0.0, 1e-3, -1e-3], np.float32)
    E, G, O = _pair(fe, oracle, synth, W, H, 1000, 20, 7, 23, K4, D4)
    for g, o in zip(G, O):
        ku = o.keys_un()
        assert g.N == o.N and np.array_equal(g.mvKeysUn.view(np.uint8), ku.view(np.uint8))
        off, idx = o.grid()
        assert np.array_equal(g.grid_offsets, off) and np.array_equal(g.grid_indices, idx)
    n, m12, cnt = fe.ORBmatcher(0.9, True).SearchForInitialization(G[0], G[1], 100)
    rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], 0.9, True, 100)
    assert n == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt)
    E.close()


def test_stream_run_back_to_back(fe, synth):
    """orbx_stream_run calls enqueued one after the other without a synchronisation (the batch runs as two halves on two lanes that
    join on the context's stream), uploads in between, runs of different sizes in a row: results must be those of a single run of
    whatever batch is resident."""
    W, H, B = 640, 480, 128
    K4, D4 = synth.camera_for(W, H)
    E = fe.ORBextractor(1000, 1.2, 8, 20, 7, W, H, max_batch=B)
    E.set_camera(K4, D4)
    S = fe.Stream(E, 0.9, True, 100)
    batches = [synth.frames(B, W, H, first_pair=300), synth.frames(B, W, H, first_pair=700)]
    single = []
    for imgs in batches:
        S.upload(imgs)
        S.run(B)
        single.append(S.download(B))
    assert not np.array_equal(single[0]["n"], single[1]["n"]) or not np.array_equal(single[0]["desc"], single[1]["desc"])

    def same(a, b):
        return all(np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)) for k in a)

    for rep in range(3):
        for j, imgs in enumerate(batches):
            S.upload(imgs)
            for _ in range(5 + rep):
                S.run(B)
            assert same(S.download(B), single[j]), (rep, j)
    # runs of different sizes in a row, and a sync in between
    S.upload(batches[0])
    S.run(B); S.run(B // 2); S.run(B); S.run(B); S.sync(); S.run(B); S.run(B)
    assert same(S.download(B), single[0])
    E.close()
