"""CPU restatement (numpy, small cases) of the row-sharded matcher's host logic, for the CPU test suite: per query the
sorted list of its best candidates (what orbx_match_rows returns for a row block) and the ordered resolve from the
gathered table (orbx_match_resolve).  Candidates and their order come from the oracle's GetFeaturesInArea; the result must
equal the oracle's SearchForInitialization (Features/ORBmatcher.cpp:11-150).  Test infrastructure only."""
import numpy as np

K, TH_LOW, HISTO = 7, 50, 30
NOKEY = 0xFFFFFFFF


def level0(F):
    return np.nonzero(~(F.keys_un()["octave"] > 0))[0]


def _candidates(F2, kp, window):
    """(positions in F2's level-0 visiting order, keypoint indices) of the query's candidates."""
    off, idx = F2.grid()
    oct2 = F2.keys_un()["octave"]
    order = idx[oct2[idx] == 0]                      # level-0 keys of F2 in visiting order
    pos = {int(i2): t for t, i2 in enumerate(order)}
    c = F2.features_in_area(float(kp["x"]), float(kp["y"]), float(window), 0, 0)
    return np.array([pos[int(i)] for i in c], np.int64), c, order


def rows(F1, F2, window, q0, q1, k=K):
    """Top lists of the queries [q0, q1): [n, 8] uint32 (k keys = dist << 20 | position, info = count | exact << 24)."""
    ku1 = F1.keys_un()
    out = np.full((max(q1 - q0, 0), 8), NOKEY, np.uint32)
    for r, i1 in enumerate(level0(F1)[q0:q1]):
        t, c, _ = _candidates(F2, ku1[i1], window)
        if len(c):
            d = np.unpackbits(F1.desc[i1][None, :] ^ F2.desc[c], axis=1).sum(1).astype(np.uint32)
            keys = np.sort((d << 20) | t.astype(np.uint32))[:k]
            out[r, :len(keys)] = keys
        out[r, 7] = len(c) | (min(len(c), k) << 24)
    return out


def resolve(F1, F2, table, nnratio, check_ori, window):
    """-> (nmatches, matches12, counters[3]) from the complete row table, with the full ordered walk as the fallback."""
    ku1, ku2 = F1.keys_un(), F2.keys_un()
    q = level0(F1)
    assert len(table) == len(q)
    m12 = np.full(F1.N, -1, np.int32)
    md = np.full(F2.N, 1 << 30, np.int64)
    m21 = np.full(F2.N, -1, np.int64)
    hist = [[] for _ in range(HISTO)]
    nmatches = bad_d = bad_r = 0
    f32 = np.float32
    for row, i1 in zip(table, q):
        ncand, nlist = int(row[7]) & 0xFFFFF, int(row[7]) >> 24
        if ncand == 0:
            continue
        t, c, order = _candidates(F2, ku1[i1], window)
        found = []
        for key in row[:nlist]:
            dist, i2 = int(key) >> 20, int(order[int(key) & 0xFFFFF])
            if md[i2] <= dist:
                continue
            found.append((dist, i2))
            if len(found) == 2:
                break
        if len(found) < 2 and nlist < ncand:   # the list ran out: walk every candidate in order
            d = np.unpackbits(F1.desc[i1][None, :] ^ F2.desc[c], axis=1).sum(1)
            keep = [(int(dd), int(tt), int(ii)) for dd, tt, ii in zip(d, t, c) if not md[ii] <= dd]
            keep.sort()
            found = [(dd, ii) for dd, tt, ii in keep[:2]]
        best = found[0][0] if found else 1 << 31
        best2 = found[1][0] if len(found) > 1 else (1 << 31) - 1
        if best > TH_LOW:
            bad_d += 1
            continue
        if f32(best) > f32(nnratio) * f32(best2):
            bad_r += 1
            continue
        i2 = found[0][1]
        if m21[i2] >= 0:
            m12[m21[i2]] = -1
            nmatches -= 1
        m12[i1] = i2; m21[i2] = i1; md[i2] = best
        nmatches += 1
        if check_ori:
            rot = f32(ku1[i1]["angle"]) - f32(ku2[i2]["angle"])
            if rot < 0:
                rot = f32(rot + f32(360.0))
            v = float(f32(rot * f32(f32(HISTO) / f32(360.0))))
            b = int(np.floor(abs(v) + 0.5) * (1 if v >= 0 else -1))   # C round(): half away from zero
            if b == HISTO:
                b = 0
            hist[b].append(i1)
    bad_o = 0
    if check_ori:
        sizes = [len(h) for h in hist]
        m1 = m2 = m3 = 0
        i1m = i2m = i3m = -1
        for i, s in enumerate(sizes):
            if s > m1:
                m3, m2, m1, i3m, i2m, i1m = m2, m1, s, i2m, i1m, i
            elif s > m2:
                m3, m2, i3m, i2m = m2, s, i2m, i
            elif s > m3:
                m3, i3m = s, i
        if f32(m2) < f32(0.1) * f32(m1):
            i2m = i3m = -1
        elif f32(m3) < f32(0.1) * f32(m1):
            i3m = -1
        for i, h in enumerate(hist):
            if i in (i1m, i2m, i3m):
                continue
            for i1 in h:          # unconditional, as the reference (:132-137): an overridden i1 counts twice
                m12[i1] = -1
                nmatches -= 1
                bad_o += 1
    return nmatches, m12, np.array([bad_d, bad_r, bad_o], np.int32)
