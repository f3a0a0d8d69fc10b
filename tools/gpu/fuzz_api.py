"""Ad-hoc fuzz of the remaining entry points on random configurations: partial batches (n < max_batch, two async steps of
different sizes in flight), edge parameters (tiny feature counts, one level, thresholds up to 255), radius queries (single and
batched, level ranges, windows outside the image), the row-split matcher with random splits, Initializer scoring with random
and degenerate hypotheses, raw Hamming best / second best (both entry points) on ragged sizes, colour ingest on odd widths."""
import sys, ctypes as C, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import test_gpu_fuzz as T
from oracle import orb_oracle as orc
from orb_slam_tracking_b200 import frontend as fe, synth, capi
from orb_slam_tracking_b200.capi import OrbxError
N, seed = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(seed)
bad = 0
ran = 0

def report(cfg, why):
    global bad
    bad += 1
    print("MISMATCH", cfg, why[:4])

for cfg in T._configs(N, seed):
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    if W * H > 500 * 400: W, H = W // 2 + 80, H // 2 + 60
    pick = idx % 4
    if pick == 0: nf = int(rng.integers(1, 12))
    elif pick == 1: nl, sf = 1, float(rng.choice([1.0, 1.2]))
    elif pick == 2: ini, mn = int(rng.choice([255, 200, 128, 127])), int(rng.choice([255, 128, 100, 5]))
    nf = min(nf, 700)
    B = int(rng.integers(6, 20)) * 2
    try:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H, max_batch=B)
    except OrbxError as e:
        continue
    R = orc.Extractor(nf, sf, nl, ini, mn)
    K4 = np.array([0.8 * W, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25], np.float32)
    D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if dist else [0, 0, 0, 0], np.float32)
    b = E.set_camera(K4, D4)
    base = synth.frames(6, W, H, first_pair=700 + 3 * idx)
    ref = []
    for im in base:
        k, d, _ = R(im); ref.append((k, d, orc.Frame(k, d, K4, D4, b)))
    why = []
    S = fe.Stream(E, 0.9, True, 100)
    # two async steps of different (partial) sizes in flight, then a third
    sizes = [2 * int(rng.integers(1, B // 2 + 1)) for _ in range(3)]
    batches = [np.ascontiguousarray(np.stack([base[(i + s) % 6] for i in range(n)])) for s, n in enumerate(sizes)]
    t0 = S.process_async(batches[0]); t1 = S.process_async(batches[1]); o0 = S.wait(t0); t2 = S.process_async(batches[2])
    outs = [o0, S.wait(t1), S.wait(t2)]
    for s, (n, out) in enumerate(zip(sizes, outs)):
        for i in range(n):
            rk, rd, _ = ref[(i + s) % 6]
            m = int(out["n"][i])
            if not (m == len(rk) and np.array_equal(out["kps"][i][:m].view(np.uint8), rk.view(np.uint8)) and np.array_equal(out["desc"][i][:m], rd)):
                why.append(("async frame", s, i, m, len(rk)))
        for p in range(n // 2):
            f1, f2 = ref[(2 * p + s) % 6][2], ref[(2 * p + 1 + s) % 6][2]
            rn, rm12, rcnt = orc.search_for_initialization(f1, f2, 0.9, True, 100)
            if not (int(out["nmatches"][p]) == rn and np.array_equal(out["matches12"][p][:f1.N], rm12) and np.array_equal(out["counters"][p], rcnt)):
                why.append(("async pair", s, p))
    # radius queries: resident frames of the last step (batched) and host frames (single)
    n = sizes[2]
    q = np.stack([rng.uniform(-60, W + 60, 40), rng.uniform(-60, H + 60, 40), rng.uniform(1, 200, 40),
                  rng.integers(-1, nl, 40), rng.integers(-1, nl + 1, 40)], axis=1).astype(np.float32)
    slots = rng.integers(0, n, 40).astype(np.int32)
    res, cnt = S.features_in_area(slots, q)
    for (x, y, r, l0, l1), sl, got in zip(q, slots, res):
        want = ref[(int(sl) + 2) % 6][2].features_in_area(float(x), float(y), float(r), int(l0), int(l1))
        if not np.array_equal(got, want): why.append(("area batched", float(x), float(y), float(r), int(l0), int(l1)))
    G = [fe.Frame(E, base[j]) for j in range(2)]
    for (x, y, r, l0, l1) in q[:10]:
        if not np.array_equal(G[1].GetFeaturesInArea(float(x), float(y), float(r), int(l0), int(l1)),
                              ref[1][2].features_in_area(float(x), float(y), float(r), int(l0), int(l1))): why.append(("area single",))
    # host-frame matcher: every path, and the row-split form with a random split
    M = fe.ORBmatcher(float(rng.choice([0.6, 0.8, 0.9])), bool(rng.integers(0, 2)))
    for win in (100, int(rng.integers(1, 60)), 5000):
        rn, rm12, rcnt = orc.search_for_initialization(ref[0][2], ref[1][2], M.mfNNratio, M.mbCheckOrientation, win)
        for path in (0, 1, 2, 3):
            E.set_option("match_path", path)
            n_, m12, cnt_ = M.SearchForInitialization(G[0], G[1], win)
            if not (n_ == rn and np.array_equal(m12, rm12) and np.array_equal(cnt_, rcnt)): why.append(("match", path, win))
        E.set_option("match_path", 0)
        try:
            _, nq = M.match_rows(G[0], G[1], win, 0, 0)
            cut = sorted(set([0, nq] + [int(c) for c in rng.integers(0, nq + 1, 2)]))
            rows = np.concatenate([M.match_rows(G[0], G[1], win, a, b_)[0] for a, b_ in zip(cut[:-1], cut[1:])] or [np.zeros((0, 8), np.uint32)])
            n_, m12, cnt_ = M.match_resolve(G[0], G[1], rows, win)
            if not (n_ == rn and np.array_equal(m12, rm12) and np.array_equal(cnt_, rcnt)): why.append(("rows", win, nq, cut))
        except OrbxError as e:
            why.append(("rows error", str(e)))
    # Initializer scoring: random, identity, zero and huge hypotheses
    rn, rm12, _ = orc.search_for_initialization(ref[0][2], ref[1][2], 0.9, True, 100)
    nh = int(rng.integers(1, 40))
    Hs = np.tile(np.eye(3, dtype=np.float32).reshape(1, 9), (nh, 1)); Hs[:, [2, 5]] += rng.uniform(-30, 30, (nh, 2)).astype(np.float32)
    Hs[:, [0, 4]] += rng.normal(0, 0.02, (nh, 2)).astype(np.float32)
    if nh > 3: Hs[1] = 0; Hs[2] = 1e30; Hs[3, 8] = 0
    Hi = Hs.copy(); Hi[:, [2, 5]] *= -1
    Fs = rng.normal(0, 1e-3, (nh, 9)).astype(np.float32)
    I = fe.Initializer(G[0], sigma=float(rng.choice([1.0, 0.5, 2.0])))
    ku1, ku2 = ref[0][2].keys_un(), ref[1][2].keys_un()
    if len(ku1) and len(ku2):
        s, inl, best = I.CheckHomography(G[1], rm12, Hs, Hi)
        rs, rinl = orc.check_homography(ku1, ku2, rm12, Hs, Hi, I.mSigma)
        fin = np.isfinite(rs)
        if not (np.array_equal(s[fin].view(np.uint32), rs[fin].view(np.uint32)) and np.array_equal(np.isnan(s), np.isnan(rs)) and best == orc.ransac_best(rs) and np.array_equal(inl[best], rinl[best])):
            why.append(("homography", nh))
        s, inl, best = I.CheckFundamental(G[1], rm12, Fs)
        rs, rinl = orc.check_fundamental(ku1, ku2, rm12, Fs, I.mSigma)
        if not (np.array_equal(s.view(np.uint32), rs.view(np.uint32)) and best == orc.ransac_best(rs) and np.array_equal(inl[best], rinl[best])):
            why.append(("fundamental", nh))
    # raw Hamming best / second best on ragged sizes (ties everywhere: few distinct descriptors)
    na, nb = int(rng.integers(1, 300)), int(rng.integers(1, 300))
    pool = rng.integers(0, 256, (int(rng.integers(2, 40)), 32), dtype=np.uint8)
    da, db = pool[rng.integers(0, len(pool), na)], pool[rng.integers(0, len(pool), nb)]
    bi, bd, sd, _ = E.hamming_best2(da, db)
    D = np.unpackbits(da[:, None, :] ^ db[None, :, :], axis=2).sum(axis=2).astype(np.int64)
    wbi = D.argmin(axis=1); wbd = D.min(axis=1)
    D2 = D.copy(); D2[np.arange(na), wbi] = 1 << 30
    wsd = D2.min(axis=1) if nb > 1 else None
    if not (np.array_equal(bi, wbi) and np.array_equal(bd, wbd) and (nb == 1 or np.array_equal(sd, wsd))): why.append(("hamming_best2", na, nb))
    # colour ingest on this width (odd widths included): BGR and RGB against the grey image the oracle's formula gives
    col = rng.integers(0, 256, (2, H, W, 3), dtype=np.uint8)
    for rgb in (False, True):
        S.upload_color(col, rgb=rgb)
        c = (col[..., ::-1] if rgb else col).astype(np.int64)
        want = ((3735 * c[..., 0] + 19235 * c[..., 1] + 9798 * c[..., 2] + 16384) >> 15).astype(np.uint8)
        S.run(2); S.sync()
        if not (np.array_equal(E.pyramid_level(0, slot=0), want[0]) and np.array_equal(E.pyramid_level(0, slot=1), want[1])): why.append(("colour", rgb))
    ran += 1
    if why: report((W, H, nf, sf, nl, ini, mn, dist, idx, B, sizes), why)
    E.close()
print("configs", N, "ran", ran, "mismatches", bad)
