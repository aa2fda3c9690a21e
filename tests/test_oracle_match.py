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
