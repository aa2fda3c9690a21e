"""CPU: the matcher oracle (oracle/orb_match.c) against an independent, deliberately naive Python restatement of
the same reference loops (src/ORBmatcher.cc:255-419, 760-903, 2076-2143; TemplatedVocabulary.h:1255-1306).
The reference has no tests or golden vectors for this path (parity unpinned, see oracle/orb_oracle.h); this keeps
the C oracle honest."""
import numpy as np
import synth, oracle_lib as O


def hamming(a, b):
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def test_descriptor_distance_vs_numpy():
    d = synth.descriptors(1, 200)
    L = O.lib()
    for i in range(0, 200, 2):
        assert L.orc_descriptor_distance(d[i].ctypes.data, d[i + 1].ctypes.data) == hamming(d[i], d[i + 1])
    assert L.orc_descriptor_distance(d[0].ctypes.data, d[0].ctypes.data) == 0
    z, o = np.zeros(32, np.uint8), np.full(32, 255, np.uint8)
    assert L.orc_descriptor_distance(z.ctypes.data, o.ctypes.data) == 256


def py_three_maxima(c):
    max1 = max2 = max3 = 0
    i1 = i2 = i3 = -1
    for i, s in enumerate(c):
        if s > max1:
            max3, max2, max1, i3, i2, i1 = max2, max1, s, i2, i1, i
        elif s > max2:
            max3, max2, i3, i2 = max2, s, i2, i
        elif s > max3:
            max3, i3 = s, i
    if max2 < np.float32(0.1) * np.float32(max1):
        i2 = i3 = -1
    elif max3 < np.float32(0.1) * np.float32(max1):
        i3 = -1
    return i1, i2, i3


def test_three_maxima():
    rng = np.random.default_rng(0)
    for _ in range(300):
        c = rng.integers(0, rng.integers(1, 40), 30)
        assert O.three_maxima(c) == py_three_maxima(list(c))
    assert O.three_maxima([0] * 30) == (-1, -1, -1)
    assert O.three_maxima([5] * 30) == (0, 1, 2)
    assert O.three_maxima([100] + [9] * 29) == (0, -1, -1)


def py_transform(v, d, levelsup):
    nid_level = v["L"] - levelsup
    node, nid, level = 0, 0, 0
    while v["child_off"][node + 1] > v["child_off"][node]:
        level += 1
        ch = v["child"][v["child_off"][node]:v["child_off"][node + 1]]
        dist = [hamming(d, v["desc"][c]) for c in ch]
        node = int(ch[int(np.argmin(dist))])  # argmin = first minimum = strict '<'
        if level == nid_level:
            nid = node
    return v["word_id"][node], v["weight"][node], nid


def test_bow_transform_vs_python():
    v = synth.vocabulary(3, 6, 3)
    vo = O.VocabArrays(v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    d = synth.descriptors(9, 120)
    for levelsup in (0, 1, 2, 3, 5):
        w, wt, nd = O.bow_transform(vo, d, levelsup)
        for i in range(len(d)):
            a, b, c = py_transform(v, d[i], levelsup)
            assert (w[i], wt[i], nd[i]) == (a, b, c)
        nid, noff, feat = O.feature_vector(nd, wt)
        fv = {}
        for i in range(len(d)):
            if wt[i] > 0:
                fv.setdefault(int(nd[i]), []).append(i)
        assert list(nid) == sorted(fv)
        for k, n in enumerate(nid):
            assert list(feat[noff[k]:noff[k + 1]]) == fv[int(n)]


def py_search_by_bow(fv1, fv2, d1, d2, a1, a2, has1, has2, nnratio, check_ori, kf_kf):
    n_out = len(d1) if kf_kf else len(d2)
    match = [-1] * n_out
    taken2 = [False] * len(d2)
    hist = [[] for _ in range(30)]
    factor = np.float32(30) / np.float32(360)  # HISTO_LENGTH / 360.0f
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if not has1[i1]:
                continue
            best1 = best2 = 256
            bi = -1
            for i2 in fv2[node]:
                if taken2[i2] or (kf_kf and not has2[i2]):
                    continue
                dist = hamming(d1[i1], d2[i2])
                if dist < best1:
                    best2, best1, bi = best1, dist, i2
                elif dist < best2:
                    best2 = dist
            ok = best1 < 50 if kf_kf else best1 <= 50
            if ok and np.float32(best1) < np.float32(nnratio) * np.float32(best2):
                taken2[bi] = True
                if kf_kf:
                    match[i1] = bi
                else:
                    match[bi] = i1
                if check_ori:
                    rot = np.float32(a1[i1]) - np.float32(a2[bi])
                    if rot < 0:
                        rot = np.float32(rot + np.float32(360))
                    x = np.float32(rot * factor)
                    b = int(np.floor(x + np.float32(0.5))) if x >= 0 else int(np.ceil(x - np.float32(0.5)))  # roundf
                    if float(x) - np.floor(float(x)) == 0.5:
                        b = int(np.floor(float(x))) + 1
                    if b == 30:
                        b = 0
                    hist[b].append(i1 if kf_kf else bi)
    if check_ori:
        keep = py_three_maxima([len(x) for x in hist])
        for b in range(30):
            if b not in keep:
                for j in hist[b]:
                    match[j] = -1
    return match


def test_search_by_bow_vs_python():
    v = synth.vocabulary(4, 5, 2)
    vo = O.VocabArrays(v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    rng = np.random.default_rng(2)
    for trial in range(4):
        n1, n2 = 160 + 10 * trial, 150
        d1 = synth.descriptors(10 + trial, n1)
        if trial >= 2:  # duplicates -> masking matters
            d1 = d1[rng.integers(0, 40, n1)]
        d2 = np.concatenate([synth.descriptors(20 + trial, 100, 0.06, d1[rng.permutation(n1)[:100]]), synth.descriptors(30 + trial, n2 - 100)])
        k1 = synth.keypoints(40 + trial, n1, dtype=O.KP_DTYPE)
        k2 = synth.keypoints(50 + trial, n2, dtype=O.KP_DTYPE)
        has1 = (rng.random(n1) < 0.8).astype(np.uint8)
        has2 = (rng.random(n2) < 0.8).astype(np.uint8)
        sides, fvs = [], []
        for d, k, has in ((d1, k1, has1), (d2, k2, has2)):
            w, wt, nd = O.bow_transform(vo, d, 1 if trial % 2 else 2)
            nid, noff, feat = O.feature_vector(nd, wt)
            sides.append(O.Side(d, k, has, None, nid, noff, feat))
            fvs.append({int(n): list(feat[noff[i]:noff[i + 1]]) for i, n in enumerate(nid)})
        for kf_kf in (False, True):
            for nnratio, ori in ((0.7, True), (0.9, False)):
                got, n = O.search_by_bow(sides[0], sides[1], nnratio, ori, kf_kf)
                want = py_search_by_bow(fvs[0], fvs[1], d1, d2, k1["angle"], k2["angle"], has1, has2, nnratio, ori, kf_kf)
                assert list(got) == want
                assert n == sum(1 for x in want if x >= 0)


def test_features_in_area_vs_bruteforce():
    keys = synth.keypoints(3, 800, 640, 480, dtype=O.KP_DTYPE)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    off, idx = O.assign_grid(keys, grid)
    assert off[-1] == len(keys) and sorted(idx) == list(range(len(keys)))
    g = O.Grid(*grid)
    rng = np.random.default_rng(1)
    out = np.empty(len(keys), np.int32)
    idxp = np.concatenate([idx, np.zeros(1, np.int32)])
    for _ in range(200):
        x, y, r = np.float32(rng.random() * 640), np.float32(rng.random() * 480), np.float32(rng.random() * 40 + 1)
        lo, hi = int(rng.integers(-1, 4)), int(rng.integers(-1, 8))
        n = O._mlib().orc_features_in_area(keys.ctypes.data, g, off.ctypes.data, idxp.ctypes.data, x, y, r, lo, hi, out.ctypes.data)
        sel = (np.abs(keys["x"] - x) < r) & (np.abs(keys["y"] - y) < r)
        if lo > 0 or hi >= 0:
            sel &= keys["octave"] >= lo
            if hi >= 0:
                sel &= keys["octave"] <= hi
        # the grid window (floor/ceil of the cell span) always covers the square, so the sets must agree
        assert sorted(out[:n]) == list(np.flatnonzero(sel))


# ---- independent (naive) restatements of the window matchers: brute-force candidate sets in grid iteration order ----
def py_features_in_area(keys, grid, x, y, r, lo, hi):
    """Frame::GetFeaturesInArea (src/Frame.cc:590-671): candidates in column-major cell order, insertion order inside a cell."""
    min_x, min_y, inv_w, inv_h = (np.float32(v) for v in grid)
    x, y, r = np.float32(x), np.float32(y), np.float32(r)
    c0 = max(0, int(np.floor(np.float32(np.float32(x - min_x) - r) * inv_w)))
    if c0 >= 64:
        return []
    c1 = min(63, int(np.ceil(np.float32(np.float32(x - min_x) + r) * inv_w)))
    if c1 < 0:
        return []
    r0 = max(0, int(np.floor(np.float32(np.float32(y - min_y) - r) * inv_h)))
    if r0 >= 48:
        return []
    r1 = min(47, int(np.ceil(np.float32(np.float32(y - min_y) + r) * inv_h)))
    if r1 < 0:
        return []
    # cell of every keypoint (AssignFeaturesToGrid, src/Frame.cc:398-422 / PosInGrid :682-698)
    px = np.round(np.float32(keys["x"] - min_x) * inv_w).astype(np.int64)
    py = np.round(np.float32(keys["y"] - min_y) * inv_h).astype(np.int64)
    out = []
    for ix in range(c0, c1 + 1):
        for iy in range(r0, r1 + 1):
            for j in np.flatnonzero((px == ix) & (py == iy)):
                o = int(keys["octave"][j])
                if lo > 0 or hi >= 0:
                    if o < lo or (hi >= 0 and o > hi):
                        continue
                if abs(np.float32(keys["x"][j] - x)) < r and abs(np.float32(keys["y"][j] - y)) < r:
                    out.append(int(j))
    return out


def py_rot_bin(a1, a2):
    rot = np.float32(a1) - np.float32(a2)
    if rot < 0:
        rot = np.float32(rot + np.float32(360.0))
    b = int(round_half_away(np.float32(rot * np.float32(30 / 360.0))))
    return 0 if b == 30 else b


def round_half_away(x):
    return np.floor(x + 0.5) if x >= 0 else np.ceil(x - 0.5)


def py_prune_hist(hist, assign_clear):
    i1, i2, i3 = py_three_maxima([len(h) for h in hist])
    removed = 0
    for i, h in enumerate(hist):
        if i in (i1, i2, i3):
            continue
        for j in h:
            removed += assign_clear(j)
    return removed


def py_window_search(pts, keys, desc, scale, grid, occupied, th, levels, accept, check_ori, level_in_area=True):
    occ = list(occupied) if occupied is not None else [0] * len(keys)
    assign = [-1] * len(keys)
    hist = [[] for _ in range(30)]
    n = 0
    for i, p in enumerate(pts):
        lvl = int(p["level"])
        lo, hi = levels(lvl)
        r = np.float32(np.float32(th) * scale[lvl])
        cand = py_features_in_area(keys, grid, p["u"], p["v"], r, lo, hi) if level_in_area else \
            [j for j in py_features_in_area(keys, grid, p["u"], p["v"], r, -1, -1)]
        best, bi = 256, -1
        for j in cand:
            if occ[j]:
                continue
            if not level_in_area and not (lo <= keys["octave"][j] <= hi):
                continue
            d = hamming(p["desc"], desc[j])
            if d < best:
                best, bi = d, j
        if best <= accept:
            assign[bi] = i; occ[bi] = 1; n += 1
            if check_ori:
                hist[py_rot_bin(p["angle"], keys["angle"][bi])].append(bi)
    if check_ori:
        def clear(j):
            assign[j] = -1
            return 1
        n -= py_prune_hist(hist, clear)
    return assign, n


def _window_scene(seed, n, nq):
    rng = np.random.default_rng(seed)
    keys = synth.keypoints(seed, n, 640, 480, dtype=O.KP_DTYPE)
    desc = synth.descriptors(seed + 1, n)
    src = rng.integers(0, n, nq)
    pts = np.zeros(nq, O.PROJ_DTYPE)
    pts["u"] = keys["x"][src] + rng.normal(0, 3, nq).astype(np.float32)
    pts["v"] = keys["y"][src] + rng.normal(0, 3, nq).astype(np.float32)
    pts["u"][:4] = [-30, 900, 100, 320]; pts["v"][:4] = [100, 100, -50, 700]
    pts["level"] = np.clip(keys["octave"][src] + rng.integers(-1, 2, nq), 0, 7)
    pts["angle"] = np.mod(keys["angle"][src] + 20 + rng.normal(0, 4, nq), 360).astype(np.float32)
    pts["desc"] = synth.descriptors(seed + 2, nq, 0.12, desc[src])
    occupied = (rng.random(n) < 0.1).astype(np.uint8)
    return keys, desc, pts, occupied


def test_search_by_projection_reloc_and_sim3_vs_python():
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    for seed, n, nq in ((5, 500, 300), (6, 900, 700)):
        keys, desc, pts, occupied = _window_scene(seed, n, nq)
        for th, orb_dist, ori in ((10.0, 100, True), (3.0, 64, True), (10.0, 100, False)):
            got, ng = O.search_by_projection_reloc(pts, keys, desc, scale, grid, occupied, th, orb_dist, ori)
            want, nw = py_window_search(pts, keys, desc, scale, grid, occupied, th, lambda l: (l - 1, l + 1), orb_dist, ori)
            assert list(got) == want and ng == nw
        for th in (10, 4):
            got, ng = O.search_by_projection_sim3(pts, keys, desc, scale, grid, occupied, th)
            want, nw = py_window_search(pts, keys, desc, scale, grid, occupied, th, lambda l: (l - 1, l), 50, False, level_in_area=False)
            assert list(got) == want and ng == nw and nw > 0


def py_search_for_initialization(keys1, desc1, prev, keys2, desc2, grid, window, nnratio, check_ori):
    n1, n2 = len(keys1), len(keys2)
    m12, m21, md = [-1] * n1, [-1] * n2, [2 ** 31 - 1] * n2
    hist = [[] for _ in range(30)]
    prev = np.array(prev, np.float32).copy()
    n = 0
    for i1 in range(n1):
        if keys1["octave"][i1] > 0:
            continue
        cand = py_features_in_area(keys2, grid, prev[i1, 0], prev[i1, 1], window, 0, 0)
        best, best2, bi = 2 ** 31 - 1, 2 ** 31 - 1, -1
        for i2 in cand:
            d = hamming(desc1[i1], desc2[i2])
            if md[i2] <= d:
                continue
            if d < best:
                best2, best, bi = best, d, i2
            elif d < best2:
                best2 = d
        if best <= 50 and np.float32(best) < np.float32(np.float32(best2) * np.float32(nnratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1; n -= 1
            m12[i1] = bi; m21[bi] = i1; md[bi] = best; n += 1
            if check_ori:
                hist[py_rot_bin(keys1["angle"][i1], keys2["angle"][bi])].append(i1)
    if check_ori:
        def clear(i1):
            if m12[i1] >= 0:
                m12[i1] = -1
                return 1
            return 0
        n -= py_prune_hist(hist, clear)
    for i1 in range(n1):
        if m12[i1] >= 0:
            prev[i1] = (keys2["x"][m12[i1]], keys2["y"][m12[i1]])
    return m12, n, prev


def init_scene(seed, n1, n2, shift=6.0, dup=0.3):
    """Two frames of level-0-heavy keypoints: frame 2 = frame 1 moved by a few pixels, descriptors perturbed; a share of
    duplicated descriptors makes several F1 keypoints compete for (and steal) the same F2 keypoint."""
    rng = np.random.default_rng(seed)
    k1 = synth.keypoints(seed, n1, 640, 480, dtype=O.KP_DTYPE)
    k1["octave"] = np.where(rng.random(n1) < 0.7, 0, k1["octave"])
    d1 = synth.descriptors(seed + 1, n1)
    ndup = int(dup * n1)
    srcd = rng.integers(0, n1, ndup); dstd = rng.integers(0, n1, ndup)
    d1[dstd] = synth.descriptors(seed + 5, ndup, 0.03, d1[srcd])
    k1["x"][dstd] = np.clip(k1["x"][srcd] + rng.normal(0, 20, ndup), 0, 639).astype(np.float32)
    k1["y"][dstd] = np.clip(k1["y"][srcd] + rng.normal(0, 20, ndup), 0, 479).astype(np.float32)
    src = rng.integers(0, n1, n2)
    k2 = k1[src].copy()
    k2["x"] = np.clip(k2["x"] + shift + rng.normal(0, 2, n2), 0, 639).astype(np.float32)
    k2["y"] = np.clip(k2["y"] + rng.normal(0, 2, n2), 0, 479).astype(np.float32)
    k2["angle"] = np.mod(k2["angle"] + 10 + rng.normal(0, 5, n2), 360).astype(np.float32)
    d2 = synth.descriptors(seed + 2, n2, 0.06, d1[src])
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    return k1, d1, prev, k2, d2


def test_search_for_initialization_vs_python():
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    stolen = 0
    for seed, n1, n2, window, nnratio, ori in ((1, 400, 380, 100, 0.9, True), (2, 600, 700, 100, 0.9, True), (3, 300, 300, 30, 0.8, False)):
        k1, d1, prev, k2, d2 = init_scene(seed, n1, n2)
        got, ng, gprev = O.search_for_initialization(k1, d1, prev, k2, d2, grid, window, nnratio, ori)
        want, nw, wprev = py_search_for_initialization(k1, d1, prev, k2, d2, grid, window, nnratio, ori)
        assert list(got) == want and ng == nw and nw > 20
        assert np.array_equal(gprev, wprev)
        assert ng == sum(1 for x in want if x >= 0)


def py_window_best(pts, keys, desc, uright, scale, grid, th, inv_sigma2, th_dist):
    bi, bd = [], []
    for p in pts:
        lvl = int(p["level"])
        r = np.float32(np.float32(th) * scale[lvl])
        best, idx = 2 ** 31 - 1, -1
        for j in py_features_in_area(keys, grid, p["u"], p["v"], r, -1, -1):
            o = int(keys["octave"][j])
            if o < lvl - 1 or o > lvl:
                continue
            if inv_sigma2 is not None:
                ex, ey = np.float32(p["u"] - keys["x"][j]), np.float32(p["v"] - keys["y"][j])
                e2 = np.float32(np.float32(ex * ex) + np.float32(ey * ey))
                lim = 5.99
                if uright is not None and uright[j] >= 0:
                    er = np.float32(p["ur"] - uright[j])
                    e2 = np.float32(e2 + np.float32(er * er)); lim = 7.8
                if float(np.float32(e2 * inv_sigma2[o])) > lim:
                    continue
            d = hamming(p["desc"], desc[j])
            if d < best:
                best, idx = d, j
        bd.append(best if idx >= 0 else -1)
        bi.append(idx if idx >= 0 and best <= th_dist else -1)
    return bi, bd


def test_window_best_vs_python():
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32)
    inv_sigma2 = np.array(p.inv_sigma2[:8], np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    for seed, n, nq in ((7, 500, 300), (8, 900, 500)):
        keys, desc, pts, _ = _window_scene(seed, n, nq)
        rng = np.random.default_rng(seed)
        uright = np.where(rng.random(n) < 0.5, keys["x"] - 6, -1).astype(np.float32)
        pts["ur"] = pts["u"] - 6 + rng.normal(0, 2, nq).astype(np.float32)
        for th, is2, ur, thd in ((3.0, inv_sigma2, uright, 50), (4.0, inv_sigma2, None, 50), (7.5, None, None, 100)):
            bi, bd = O.window_best(pts, keys, desc, ur, scale, grid, th, is2, thd)
            wi, wd = py_window_best(pts, keys, desc, ur, scale, grid, th, is2, thd)
            assert list(bi) == wi and list(bd) == wd
            assert sum(1 for x in wi if x >= 0) > 20


def test_distinctive_descriptors_vs_numpy():
    rng = np.random.default_rng(3)
    sizes = [1, 2, 3, 7, 16, 33, 64, 100, 0, 5]
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = synth.descriptors(1, len(sizes))
    desc = np.concatenate([synth.descriptors(10 + k, s, 0.15, np.repeat(base[k:k + 1], max(s, 1), 0)) for k, s in enumerate(sizes)])
    got = O.distinctive_descriptors(desc, seg)
    for k, s in enumerate(sizes):
        if s == 0:
            assert got[k] == -1
            continue
        D = desc[seg[k]:seg[k + 1]]
        dist = np.unpackbits(D[:, None, :] ^ D[None, :, :], axis=2).sum(2)
        med = np.sort(dist, axis=1)[:, int(0.5 * (s - 1))]
        assert got[k] == int(np.argmin(med))


def py_bow_vector(word, ww):
    v = {}
    for w in word:
        w = int(w)
        if ww[w] > 0:
            v[w] = v[w] + ww[w] if w in v else ww[w]        # python floats are IEEE doubles, added in feature order
    ids = sorted(v)
    norm = 0.0
    for i in ids:
        norm += abs(v[i])
    return ids, [v[i] / norm if norm > 0 else v[i] for i in ids]


def py_score_l1(i1, v1, i2, v2):
    d2 = dict(zip(i2, v2))
    s = 0.0
    for i, a in zip(i1, v1):
        if i in d2:
            s += abs(a - d2[i]) - abs(a) - abs(d2[i])
    return -s / 2.0


def bow_scene(seed, nwords, nfeat):
    rng = np.random.default_rng(seed)
    ww = (rng.random(nwords) * 3 + 0.01)
    ww[rng.random(nwords) < 0.05] = 0.0                      # stopped words
    hot = rng.integers(0, nwords, max(nwords // 8, 1))
    word = np.where(rng.random(nfeat) < 0.6, hot[rng.integers(0, len(hot), nfeat)], rng.integers(0, nwords, nfeat)).astype(np.int32)
    return ww, word


def test_bow_vector_and_score_vs_python():
    for seed, nwords, nfeat in ((1, 1000, 1500), (2, 100000, 4000), (3, 10, 300), (4, 50, 1)):
        ww, word = bow_scene(seed, nwords, nfeat)
        ids, vals = O.bow_vector(word, ww)
        wi, wv = py_bow_vector(word, ww)
        assert list(ids) == wi and np.array_equal(vals.view(np.uint64), np.array(wv, np.float64).view(np.uint64))
        if len(wv):
            assert abs(sum(wv) - 1.0) < 1e-9
        _, word2 = bow_scene(seed + 100, nwords, nfeat)
        word2[: nfeat // 2] = word[: nfeat // 2]
        ids2, vals2 = O.bow_vector(word2, ww)
        got = O.bow_score_l1(ids, vals, ids2, vals2)
        want = py_score_l1(list(ids), list(vals), list(ids2), list(vals2))
        assert np.float64(got).view(np.uint64) == np.float64(want).view(np.uint64)
        assert O.bow_score_l1(ids, vals, ids, vals) > 0.999 or len(ids) == 0
