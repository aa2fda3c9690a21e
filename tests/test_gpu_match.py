"""GPU parity of the Hamming side (through the C ABI) vs the CPU oracle (oracle/orb_match.c):
DBoW2 transform, both SearchByBoW overloads, SearchForTriangulation, the window (projection) search,
AssignFeaturesToGrid and ComputeStereoMatches.  Every output index must be identical."""
import os
import numpy as np
import pytest
import synth, oracle_lib as O
from gpu_util import pkg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    return pkg()


@pytest.fixture(scope="module")
def ctx(P):
    return P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)


from matcher_scene import make_pair


def sides(P, voc_o, voc_g, d1, k1, d2, k2, levelsup, seed, p_has=0.7, stereo=False):
    rng = np.random.default_rng(seed)
    out_o, out_g = [], []
    for d, k in ((d1, k1), (d2, k2)):
        w, wt, nd = O.bow_transform(voc_o, d, levelsup)
        gw, gwt, gnd = voc_g.transform(d, levelsup)
        assert np.array_equal(w, gw) and np.array_equal(nd, gnd) and np.array_equal(wt.view(np.uint32), gwt.view(np.uint32))
        nid, noff, feat = O.feature_vector(nd, wt)
        gnid, gnoff, gfeat = P.feature_vector(gnd, gwt)
        assert np.array_equal(nid, gnid) and np.array_equal(noff, gnoff) and np.array_equal(feat, gfeat)
        has = (rng.random(len(d)) < p_has).astype(np.uint8)
        ur = None
        if stereo:
            ur = np.where(rng.random(len(d)) < 0.5, k["x"] - 5, -1).astype(np.float32)
        out_o.append(O.Side(d, k, has, ur, nid, noff, feat))
        out_g.append(P.FrameView(d, k, has, ur, nid, noff, feat))
    return out_o, out_g


@pytest.fixture(scope="module")
def vocs(P, ctx):
    v = synth.vocabulary(0, 10, 3)
    vo = O.VocabArrays(v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    vg = P.Vocabulary(ctx, v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    return v, vo, vg


def test_vocab_text_loader_and_transform(P, ctx, vocs, tmp_path):
    v, vo, vg = vocs
    path = tmp_path / "voc.txt"
    path.write_text(synth.vocabulary_text(v))
    vt = P.Vocabulary(ctx, path=str(path))
    assert vt.info() == vg.info() == (10, 3, 1111, 1000)
    d = synth.descriptors(5, 3000)
    for levelsup in (0, 1, 2, 3, 4):
        a = O.bow_transform(vo, d, levelsup)
        b = vg.transform(d, levelsup)
        c = vt.transform(d, levelsup)
        for x, y, z in zip(a, b, c):
            assert np.array_equal(x.view(np.uint32), y.view(np.uint32)) and np.array_equal(y.view(np.uint32), z.view(np.uint32))


def saved_text(v):
    """What TemplatedVocabulary::saveToTextFile (TemplatedVocabulary.h:1492-1517) writes for this tree: the header with two
    blanks, every descriptor byte followed by a blank, one more blank, the weight as ostream prints a double (%g)."""
    out = [f"{v['k']} {v['L']}  0 0\n"]
    for i in range(1, len(v["parent"])):
        out.append(f"{v['parent'][i]} {v['leaf'][i]} " + "".join(f"{int(b)} " for b in v["desc"][i]) + " %g\n" % float(v["weight"][i]))
    return "".join(out)


def test_vocab_save_text_round_trip(P, ctx, vocs, tmp_path):
    """orbx_vocab_save_text writes the file the reference's saveToTextFile writes (compared with oracle/_ref where it is here,
    else with the format restated above), and loading it back gives the same vocabulary."""
    import os, ref_lib as R
    v, _, _ = vocs
    src = tmp_path / "voc.txt"
    src.write_text(synth.vocabulary_text(v).rstrip("\n"))   # the reference's loader turns a trailing newline into a phantom node
    vt = P.Vocabulary(ctx, path=str(src))
    out = tmp_path / "saved.txt"
    vt.save_text(str(out))
    got = out.read_text()
    if os.path.exists(R.SO):
        L = R.lib()
        h = L.ref_vocab_load_text(str(src).encode())
        assert h
        L.ref_vocab_save_text(h, str(tmp_path / "ref.txt").encode())
        L.ref_vocab_free(h)
        assert got == (tmp_path / "ref.txt").read_text()
    else:
        assert got == saved_text(v)
    again = P.Vocabulary(ctx, path=str(out))
    d = synth.descriptors(9, 500)
    assert again.info() == vt.info()
    for a, b in zip(vt.transform(d, 2), again.transform(d, 2)):
        assert np.array_equal(a, b) or np.allclose(a, b, rtol=1e-5)          # weights went through 6 significant digits


@pytest.fixture(scope="module")
def vocs6(P, ctx):
    """configs[3]'s vocabulary shape: k = 10, L = 6 (1 111 111 nodes = 35 MB of node descriptors, 10^6 words)."""
    v = synth.vocabulary_bfs(0, 10, 6)
    vo = O.VocabArrays(v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    vg = P.Vocabulary(ctx, v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    return v, vo, vg


def _near_leaves(v, seed, n):
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(v["leaf"] > 0)
    near = synth.descriptors(seed, n * 3 // 4, 0.06, v["desc"][rng.choice(leaves, n * 3 // 4)])
    return np.ascontiguousarray(np.concatenate([near, synth.descriptors(seed + 1, n - len(near))]))


def test_orbvoc_sized_vocabulary_transform_and_text_loader(P, ctx, vocs6, tmp_path):
    """The 6-level descent over 1.1 M nodes: GPU == oracle at levelsup 0 / 4 / 6 (6 = one node = brute-force SearchByBoW),
    from the flattened arrays and from the 170 MB ORBvoc-format text file (which DBoW2's own loader reads to the same tree:
    tests/test_oracle_vs_ref.py); BowVector of the 10 k features bit-equal as doubles."""
    v, vo, vg = vocs6
    assert vg.info() == (10, 6, 1111111, 1000000)
    d = _near_leaves(v, 7, 10000)
    want = {ls: O.bow_transform(vo, d, ls) for ls in (0, 4, 6)}
    for ls, w in want.items():
        got = vg.transform(d, ls)
        for x, y in zip(w, got):
            assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), ls
    assert len(np.unique(want[4][2])) <= 100 and len(np.unique(want[4][2])) > 50 and np.all(want[6][2] == 0)
    assert len(np.unique(want[0][0])) > 5000                                   # the descents really spread over the leaves
    path = tmp_path / "voc6.txt"
    path.write_bytes(synth.vocabulary_text_fixed(v))
    vt = P.Vocabulary(ctx, path=str(path))
    path.unlink()
    assert vt.info() == vg.info() and vt.modes() == (0, 0)
    for ls, w in want.items():
        got = vt.transform(d, ls)
        for x, y in zip(w, got):
            assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), ("text", ls)
    ww = np.zeros(1000000, np.float64)
    ww[v["word_id"][v["leaf"] > 0]] = v["weight"][v["leaf"] > 0].astype(np.float64)
    ids, vals = O.bow_vector(want[4][0], ww)
    gi, gv = vt.bow_vector(want[4][0])
    assert np.array_equal(ids, gi) and np.array_equal(vals.view(np.uint64), gv.view(np.uint64))


@pytest.mark.parametrize("levelsup", [4, 6])
def test_config4_matchers_through_the_orbvoc_sized_vocabulary(P, ctx, vocs6, levelsup):
    """configs[3]: 10 k x 10 k descriptors per keyframe pair, FeatureVectors from the k = 10, L = 6 tree at levelsup 4 (BoW
    mode, <= 100 nodes) and 6 (one node: brute force): SearchByBoW x2 and SearchForTriangulation == oracle."""
    v, vo, vg = vocs6
    n = 10000
    d1 = _near_leaves(v, 31, n)
    rng = np.random.default_rng(32)
    perm = rng.permutation(n)
    d2 = np.concatenate([synth.descriptors(33, 7000, 0.05, d1[perm[:7000]]), _near_leaves(v, 34, 3000)])
    k1, k2 = synth.keypoints(35, n, dtype=P.KP_DTYPE), synth.keypoints(36, n, dtype=P.KP_DTYPE)
    k2["angle"][:7000] = np.mod(k1["angle"][perm[:7000]] - np.float32(30) + rng.normal(0, 3, 7000).astype(np.float32), np.float32(360))
    k2["x"][:7000] = k1["x"][perm[:7000]] + rng.normal(0, 2, 7000).astype(np.float32)
    k2["y"][:7000] = k1["y"][perm[:7000]] + rng.normal(0, 2, 7000).astype(np.float32)
    k2["octave"][:7000] = k1["octave"][perm[:7000]]
    so, sg = sides(P, vo, vg, d1, k1, d2, k2, levelsup, 37, p_has=0.6, stereo=True)
    assert (len(so[0].node_id) == 1) == (levelsup == 6)
    for kf_kf in (False, True):
        want, nw = O.search_by_bow(so[0], so[1], 0.7, True, kf_kf)
        got, ng = P.ORBmatcher(ctx, 0.7, True).SearchByBoW(sg[0], sg[1], kf_kf)
        assert np.array_equal(want, got) and nw == ng and nw > 500, (kf_kf, nw, ng)
    p = O.params(1000)
    F12 = (np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + np.float32(1e-4) * rng.normal(size=(3, 3)).astype(np.float32))
    sc16 = np.zeros(16, np.float32); sc16[:8] = p.scale[:8]
    sg16 = np.zeros(16, np.float32); sg16[:8] = np.array(p.sigma2[:8], np.float32) * 9
    want, nw = O.search_for_triangulation(so[0], so[1], F12, 300.0, 200.0, sc16, sg16, False, True)
    got, ng = P.ORBmatcher(ctx, 0.6, True).SearchForTriangulation(sg[0], sg[1], F12, (300.0, 200.0), sc16, sg16, False)
    assert np.array_equal(want, got) and nw == ng and nw > 100, (nw, ng)


def test_vocabulary_modes_other_than_l1_tfidf_are_refused(P, ctx, tmp_path):
    """A text header with IDF / BINARY weighting or a non-L1 scoring: transform still works (the descent does not depend on
    them), the BowVector entry point refuses instead of accumulating the wrong way (TemplatedVocabulary.h:1166-1221)."""
    v = synth.vocabulary(0, 10, 3)
    body = synth.vocabulary_text(v).split("\n", 1)[1]
    d = synth.descriptors(5, 200)
    for scoring, weighting, ok in ((0, 0, True), (0, 1, True), (0, 2, False), (0, 3, False), (1, 0, False), (5, 0, False)):
        path = tmp_path / f"voc_{scoring}_{weighting}.txt"
        path.write_text(f"10 3 {scoring} {weighting}\n" + body)
        vt = P.Vocabulary(ctx, path=str(path))
        assert vt.modes() == (scoring, weighting)
        word, _, _ = vt.transform(d, 2)
        if ok:
            assert len(vt.bow_vector(word)[0]) > 0
        else:
            with pytest.raises(P.OrbxError, match="unsupported"):
                vt.bow_vector(word)
    bad = tmp_path / "bad.txt"
    bad.write_text("10 3 9 0\n" + body)                                   # the reference's own header gate (:1409)
    with pytest.raises(P.OrbxError):
        P.Vocabulary(ctx, path=str(bad))


CASES = [(1, 1500, 1400, 0.08, 0.0, 1), (2, 2000, 2000, 0.12, 0.0, 2), (3, 1200, 900, 0.05, 0.6, 1), (4, 700, 1800, 0.10, 0.3, 3),
         (5, 3000, 3000, 0.10, 0.0, 4), (6, 40, 35, 0.05, 0.0, 2), (7, 2500, 2500, 0.02, 0.9, 3),
         (8, 10000, 10000, 0.10, 0.0, 4), (9, 10000, 10000, 0.10, 0.0, 1)]   # config 4: 10 k x 10 k, brute force (one node) and BoW


@pytest.mark.parametrize("seed,n1,n2,flip,dup,levelsup", CASES)
@pytest.mark.parametrize("kf_kf", [False, True])
def test_search_by_bow_matches_oracle(P, ctx, vocs, seed, n1, n2, flip, dup, levelsup, kf_kf):
    _, vo, vg = vocs
    d1, k1, d2, k2 = make_pair(seed, n1, n2, flip, dup, KP=P.KP_DTYPE)
    so, sg = sides(P, vo, vg, d1, k1, d2, k2, levelsup, seed)
    for nnratio, ori in ((0.7, True), (0.9, False), (0.75, True)):
        want, nw = O.search_by_bow(so[0], so[1], nnratio, ori, kf_kf)
        got, ng = P.ORBmatcher(ctx, nnratio, ori).SearchByBoW(sg[0], sg[1], kf_kf)
        assert np.array_equal(want, got), (int((want != got).sum()), nw, ng)
        assert nw == ng
    assert nw > 0 or n1 < 100


@pytest.mark.parametrize("seed,n1,n2,flip,dup,levelsup", CASES)
@pytest.mark.parametrize("only_stereo", [False, True])
def test_search_for_triangulation_matches_oracle(P, ctx, vocs, seed, n1, n2, flip, dup, levelsup, only_stereo):
    _, vo, vg = vocs
    d1, k1, d2, k2 = make_pair(seed + 100, n1, n2, flip, dup, KP=P.KP_DTYPE)
    so, sg = sides(P, vo, vg, d1, k1, d2, k2, levelsup, seed, p_has=0.4, stereo=True)
    # a fundamental matrix for a near-pure horizontal translation: epipolar lines ~ horizontal through kp1.y
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + np.float32(1e-4) * np.random.default_rng(seed).normal(size=(3, 3)).astype(np.float32)
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32)
    sigma2 = np.array(p.sigma2[:8], np.float32)
    sc16 = np.zeros(16, np.float32); sc16[:8] = scale
    sg16 = np.zeros(16, np.float32); sg16[:8] = sigma2 * 9  # loose enough that the +-2 px jitter passes sometimes
    want, nw = O.search_for_triangulation(so[0], so[1], F12, 300.0, 200.0, sc16, sg16, only_stereo, True)
    got, ng = P.ORBmatcher(ctx, 0.6, True).SearchForTriangulation(sg[0], sg[1], F12, (300.0, 200.0), sc16, sg16, only_stereo)
    assert np.array_equal(want, got), (int((want != got).sum()), nw, ng)
    assert nw == ng


def test_assign_grid_matches_oracle(P, ctx):
    for seed, n in ((1, 1000), (2, 4000), (3, 0), (4, 17)):
        k = synth.keypoints(seed, n, 640, 480, dtype=P.KP_DTYPE)
        if n:
            k["x"][:3] = [-5, 700, 320]  # outside the grid -> dropped
        grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
        off, idx = O.assign_grid(k, grid)
        goff, gidx = P.assign_features_to_grid(ctx, k, grid)
        assert np.array_equal(off, goff) and np.array_equal(idx, gidx)


def proj_scene(P, seed, n, nq):
    import matcher_scene
    return matcher_scene.proj_scene(seed, n, nq)


@pytest.mark.parametrize("seed,n,nq", [(1, 1000, 600), (2, 2000, 2500), (3, 300, 50), (4, 4000, 4000)])
def test_search_by_projection_matches_oracle(P, ctx, seed, n, nq):
    keys, desc, uright, occupied, pts = proj_scene(P, seed, n, nq)
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    fv = P.FrameView(desc, keys, None, uright)
    # temporal points (Observations() == 0, Tracking::UpdateLastFrame) do not hide the keypoint they take (:136-138,
    # :1840-1842): none, a third of the queries, every query; the last case makes the greedy loop order-free except for
    # the "last one stays" rule.  Many queries aim at the same keypoints (nq >= n in two cases), so re-takes do happen.
    rng = np.random.default_rng(seed + 77)
    masks = [None, (rng.random(nq) < 0.35).astype(np.uint8), np.ones(nq, np.uint8)]
    # (a15) SearchByProjection(Frame, MapPoints, th): src/ORBmatcher.cc:78-191
    for (th, nnratio), tr in zip(((1.0, 0.8), (3.0, 0.8), (5.0, 0.6), (5.0, 0.9)), masks + [masks[1]]):
        want, nw = O.search_by_projection(pts, keys, desc, uright, scale, grid, occupied, th, nnratio=nnratio, transient=tr)
        q = np.zeros(nq, P.PROJ_QUERY_DTYPE)
        if tr is not None:
            q["flags"] = tr                                                     # ORBX_PROJ_QUERY_TRANSIENT
        r = np.where(pts["view_cos"] > np.float32(0.998), np.float32(2.5), np.float32(4.0)).astype(np.float32)
        if th != 1.0:
            r = (r * np.float32(th)).astype(np.float32)
        q["u"], q["v"], q["ur"], q["angle"], q["desc"] = pts["u"], pts["v"], pts["ur"], pts["angle"], pts["desc"]
        q["radius"] = (r * scale[pts["level"]]).astype(np.float32)
        q["min_level"], q["max_level"] = pts["level"] - 1, pts["level"]
        got, ng = P.ORBmatcher(ctx, nnratio, True).SearchByProjection(fv, q, grid, 100, occupied, ratio_same_level=True, check_ori=False)
        assert np.array_equal(want, got), (th, int((want != got).sum()), nw, ng)
        assert nw == ng
    # (a16) SearchByProjection(Frame, Frame, th, mono): src/ORBmatcher.cc:1741-1910
    for (th, motion), tr in zip(((15.0, 0), (7.0, 1), (15.0, 2), (15.0, 0), (25.0, 2)), masks + [masks[1], masks[2]]):
        want, nw = O.search_by_projection(pts, keys, desc, uright, scale, grid, occupied, th, motion=motion, check_ori=True, transient=tr)
        q = np.zeros(nq, P.PROJ_QUERY_DTYPE)
        if tr is not None:
            q["flags"] = tr
        q["u"], q["v"], q["ur"], q["angle"], q["desc"] = pts["u"], pts["v"], pts["ur"], pts["angle"], pts["desc"]
        q["radius"] = (np.float32(th) * scale[pts["level"]]).astype(np.float32)
        if motion == 1:
            q["min_level"], q["max_level"] = pts["level"], -1
        elif motion == 2:
            q["min_level"], q["max_level"] = 0, pts["level"]
        else:
            q["min_level"], q["max_level"] = pts["level"] - 1, pts["level"] + 1
        got, ng = P.ORBmatcher(ctx, 0.9, True).SearchByProjection(fv, q, grid, 100, occupied)
        assert np.array_equal(want, got), (th, motion, int((want != got).sum()), nw, ng)
        assert nw == ng and nw > 0


@pytest.mark.parametrize("seed,n,nq", [(11, 1000, 600), (12, 3000, 3500), (13, 200, 40)])
def test_search_by_projection_reloc_and_sim3_match_oracle(P, ctx, seed, n, nq):
    """(a17) SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, ORBdist) src/ORBmatcher.cc:1922-2065 and
    (a18) SearchByProjection(KeyFrame, Scw, vpPoints, vpMatched, th) :431-560 through the shared window matcher."""
    keys, desc, uright, occupied, pts = proj_scene(P, seed, n, nq)
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    fv = P.FrameView(desc, keys, None, uright)

    def queries(th):
        q = np.zeros(nq, P.PROJ_QUERY_DTYPE)
        q["u"], q["v"], q["angle"], q["desc"] = pts["u"], pts["v"], pts["angle"], pts["desc"]
        q["radius"] = (np.float32(th) * scale[pts["level"]]).astype(np.float32)
        return q
    for th, orb_dist, ori in ((10.0, 100, True), (3.0, 64, True), (10.0, 100, False)):
        want, nw = O.search_by_projection_reloc(pts, keys, desc, scale, grid, occupied, th, orb_dist, ori)
        q = queries(th)
        q["min_level"], q["max_level"] = pts["level"] - 1, pts["level"] + 1
        got, ng = P.ORBmatcher(ctx, 0.9, ori).SearchByProjection(fv, q, grid, orb_dist, occupied, check_uright=False)
        assert np.array_equal(want, got), (th, orb_dist, int((want != got).sum()), nw, ng)
        assert nw == ng and nw > 0
    for th in (10, 4):
        want, nw = O.search_by_projection_sim3(pts, keys, desc, scale, grid, occupied, th)
        q = queries(th)
        q["min_level"], q["max_level"] = pts["level"] - 1, pts["level"]
        got, ng = P.ORBmatcher(ctx, 0.9, False).SearchByProjection(fv, q, grid, 50, occupied, check_uright=False)
        assert np.array_equal(want, got), (th, int((want != got).sum()), nw, ng)
        assert nw == ng and nw > 0


@pytest.mark.parametrize("seed,n1,n2,window,nnratio,ori,dup", [(1, 400, 380, 100, 0.9, True, 0.3), (2, 1000, 1100, 100, 0.9, True, 0.3),
                                                              (3, 300, 300, 30, 0.8, False, 0.6), (4, 3000, 3000, 100, 0.9, True, 0.8),
                                                              (5, 50, 0, 100, 0.9, True, 0.0), (6, 2000, 2100, 200, 0.95, True, 0.9)])
def test_search_for_initialization_matches_oracle(P, ctx, seed, n1, n2, window, nnratio, ori, dup):
    """SearchForInitialization src/ORBmatcher.cc:580-748 (distance-masked candidates, stolen matches, histogram quirk)."""
    from test_oracle_match import init_scene
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    k1, d1, prev, k2, d2 = init_scene(seed, n1, max(n2, 1), dup=dup)
    k2, d2 = k2[:n2], d2[:n2]
    want, nw, wprev = O.search_for_initialization(k1, d1, prev, k2, d2, grid, window, nnratio, ori)
    got, ng, gprev = P.ORBmatcher(ctx, nnratio, ori).SearchForInitialization(P.FrameView(d1, k1), P.FrameView(d2, k2), prev, grid, window)
    assert np.array_equal(want, got), (int((want != got).sum()), nw, ng)
    assert nw == ng and (n2 == 0 or nw > 10)
    assert np.array_equal(wprev.view(np.uint32), gprev.view(np.uint32))


@pytest.mark.parametrize("seed,n,nq", [(21, 1000, 800), (22, 4000, 5000), (23, 100, 30)])
def test_window_best_matches_oracle(P, ctx, seed, n, nq):
    """Candidate loops of Fuse x2 (src/ORBmatcher.cc:1139-1300 with the reprojection gate, :1310-1458) and SearchBySim3
    (:1472-1723): independent best-only window searches."""
    keys, desc, uright, occupied, pts = proj_scene(P, seed, n, nq)
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32)
    inv_sigma2 = np.array(p.inv_sigma2[:8], np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    uright = np.where(uright > 0, keys["x"] - 6, -1).astype(np.float32)
    pts["ur"] = pts["u"] - 6 + np.random.default_rng(seed).normal(0, 2, nq).astype(np.float32)
    for th, is2, ur, thd in ((3.0, inv_sigma2, uright, 50), (4.0, inv_sigma2, None, 50), (7.5, None, None, 100), (3.0, None, uright, 50)):
        wi, wd = O.window_best(pts, keys, desc, ur, scale, grid, th, is2, thd)
        q = np.zeros(nq, P.PROJ_QUERY_DTYPE)
        q["u"], q["v"], q["ur"], q["desc"] = pts["u"], pts["v"], pts["ur"], pts["desc"]
        q["radius"] = (np.float32(th) * scale[pts["level"]]).astype(np.float32)
        q["min_level"], q["max_level"] = pts["level"] - 1, pts["level"]
        gi, gd = P.window_best(ctx, P.FrameView(desc, keys, None, ur), q, grid, thd, is2)
        assert np.array_equal(wd, gd), (th, int((wd != gd).sum()))
        assert np.array_equal(wi, gi), (th, int((wi != gi).sum()))
        assert (wi >= 0).sum() > 5


def test_bow_vector_and_l1_score_match_oracle(P, ctx, vocs, tmp_path):
    """BowVector of transform (addWeight in feature order + L1 normalisation, doubles) and L1Scoring::score against many
    database vectors: bit-equal doubles.  Also the text loader's exact double weights."""
    from test_oracle_match import bow_scene
    v, vo, vg = vocs
    nwords = vg.info()[3]
    rng = np.random.default_rng(9)
    ww = rng.random(nwords) * 3 + 0.01
    ww[rng.random(nwords) < 0.05] = 0.0
    vg.set_word_weights(ww)
    vecs_o, vecs_g = [], []
    for seed, nfeat in ((1, 1500), (2, 4000), (3, 1), (4, 10000), (5, 300), (6, 16384)):
        word = np.random.default_rng(seed).integers(0, nwords, nfeat).astype(np.int32)
        word[: nfeat // 3] = word[0]
        io, vo_ = O.bow_vector(word, ww)
        ig, vg_ = vg.bow_vector(word)
        assert np.array_equal(io, ig) and np.array_equal(vo_.view(np.uint64), vg_.view(np.uint64)), (seed, nfeat)
        vecs_o.append((io, vo_)); vecs_g.append((ig, vg_))
    with pytest.raises(P.OrbxError):
        vg.bow_vector(np.zeros(20000, np.int32))                      # more than 16384 features per call
    want = np.array([O.bow_score_l1(vecs_o[0][0], vecs_o[0][1], i, x) for i, x in vecs_o])
    got = P.bow_score_l1(ctx, vecs_g[0][0], vecs_g[0][1], vecs_g)
    assert np.array_equal(want.view(np.uint64), got.view(np.uint64))
    assert got[0] > 0.999
    vg.set_word_weights(v["weight"][v["leaf"] > 0].astype(np.float64))       # back to the fixture's weights
    # descriptors -> words -> vector through the whole chain, with a text vocabulary whose weights need all 17 digits
    v2 = dict(v)
    w64 = v["weight"].astype(np.float64)
    w64[v["leaf"] > 0] = ww
    v2["weight"] = w64
    path = tmp_path / "voc64.txt"
    path.write_text(synth.vocabulary_text(v2))
    vt = P.Vocabulary(ctx, path=str(path))
    d = synth.descriptors(77, 2000)
    w, _, _ = vt.transform(d, 4)
    ig, vg_ = vt.bow_vector(w)
    io, vo_ = O.bow_vector(w, ww)
    assert len(io) > 100 and np.array_equal(io, ig) and np.array_equal(vo_.view(np.uint64), vg_.view(np.uint64))


def test_distinctive_descriptors_match_oracle(P, ctx):
    """MapPoint::ComputeDistinctiveDescriptors src/MapPoint.cc:366-448, many map points per call."""
    rng = np.random.default_rng(5)
    sizes = [1, 2, 3, 7, 16, 33, 64, 100, 0, 5, 31, 32, 257, 1024] + list(rng.integers(1, 60, 300))
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = synth.descriptors(1, len(sizes))
    desc = np.concatenate([synth.descriptors(10 + k, s, 0.15, np.repeat(base[k:k + 1], max(s, 1), 0)) for k, s in enumerate(sizes)])
    want = O.distinctive_descriptors(desc, seg)
    got = P.distinctive_descriptors(ctx, desc, seg)
    assert np.array_equal(want, got), int((want != got).sum())
    with pytest.raises(P.OrbxError):
        P.distinctive_descriptors(ctx, synth.descriptors(2, 1025), np.array([0, 1025], np.int32))


@pytest.mark.parametrize("w,h,nf,seed", [(752, 480, 1200, 3), (1241, 376, 2000, 4)])
def test_stereo_matches_oracle(P, w, h, nf, seed):
    left, right = synth.stereo_pair(seed, w, h)
    exl = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h)
    exr = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h)
    kl, dl = exl(left)
    kr, dr = exr(right)
    p = O.params(nf)
    okl, odl, pyl = O.extract(p, left, want_pyr=True)
    okr, odr, pyr = O.extract(p, right, want_pyr=True)
    assert np.array_equal(kl, okl) and np.array_equal(kr, okr)
    sizes = O.level_sizes(p, w, h)
    bf, b = 386.1448, 386.1448 / 718.856  # Examples/Stereo/KITTI00-02.yaml: Camera.bf, fx -> mb = mbf / fx
    want_u, want_z, nv = O.stereo_matches(okl, odl, okr, odr, pyl, pyr, sizes, np.array(p.scale[:8], np.float32),
                                          np.array(p.inv_scale[:8], np.float32), bf, b)
    got_u, got_z = P.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, b)
    assert np.array_equal(want_u.view(np.uint32), got_u.view(np.uint32)), int((want_u != got_u).sum())
    assert np.array_equal(want_z.view(np.uint32), got_z.view(np.uint32))
    assert (got_u >= 0).sum() > 100, (int((got_u >= 0).sum()), nv)


@pytest.mark.parametrize("w,h,nf,pairs", [(752, 480, 1200, 64), (1241, 376, 2000, 24), (752, 480, 1200, 512)])   # 512 = configs[2] at full size
def test_stereo_batch_matches_oracle_and_device_chain(P, w, h, nf, pairs):
    """configs[2] / configs[1] as batches: left and right extract_batch, then ONE orbx_stereo_match_batch, against
    orc_extract + orc_stereo_matches pair by pair (mvuRight / mvDepth as raw bits); then the same chain with everything left
    on the device (counts included) must give the same bits."""
    import torch
    bf, b = 386.1448, 386.1448 / 718.856
    prs = [synth.stereo_pair(100 + i, w, h) for i in range(min(pairs, 6))]            # six distinct pairs, tiled
    left = np.stack([prs[i % len(prs)][0] for i in range(pairs)]); right = np.stack([prs[i % len(prs)][1] for i in range(pairs)])
    exl = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=pairs)
    exr = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=pairs)
    cap = exl.max_keypoints()
    dev = torch.device("cuda", 0)
    dl, dr = torch.from_numpy(left).to(dev), torch.from_numpy(right).to(dev)
    mk = lambda: (torch.empty((pairs, cap, 28), dtype=torch.uint8, device=dev), torch.empty((pairs, cap, 32), dtype=torch.uint8, device=dev),
                  torch.zeros(pairs, dtype=torch.int32, device=dev))
    kl, xl, nl = mk(); kr, xr, nr = mk()
    exl.extract_batch_device(dl, kl, xl, nl); exr.extract_batch_device(dr, kr, xr, nr)
    exl.sync(); exr.sync()
    ur_d, z_d, nm_d = P.compute_stereo_matches_batch(exl, exr, kl, xl, nl, kr, xr, nr, bf, b, device=True)
    exl.sync()
    hk = lambda t: t.cpu().numpy().view(P.KP_DTYPE).reshape(pairs, cap)
    ur_h, z_h, nm_h = P.compute_stereo_matches_batch(exl, exr, hk(kl), xl.cpu().numpy(), nl.cpu().numpy(), hk(kr), xr.cpu().numpy(),
                                                     nr.cpu().numpy(), bf, b)
    p = O.params(nf)
    sizes = O.level_sizes(p, w, h)
    want = {}
    for i in range(len(prs)):
        okl, odl, pyl = O.extract(p, prs[i][0], want_pyr=True)
        okr, odr, pyr = O.extract(p, prs[i][1], want_pyr=True)
        want[i] = O.stereo_matches(okl, odl, okr, odr, pyl, pyr, sizes, np.array(p.scale[:8], np.float32), np.array(p.inv_scale[:8], np.float32), bf, b)
    nl_h = nl.cpu().numpy()
    for q in range(pairs):
        wu, wz, _ = want[q % len(prs)]
        n = int(nl_h[q])
        assert n == len(wu)
        for got_u, got_z, nm in ((ur_h, z_h, nm_h), (ur_d.cpu().numpy(), z_d.cpu().numpy(), nm_d)):
            assert np.array_equal(got_u[q, :n].view(np.uint32), wu.view(np.uint32)), q
            assert np.array_equal(got_z[q, :n].view(np.uint32), wz.view(np.uint32)), q
            assert nm[q] == int((wz > 0).sum()) and nm[q] > 100
    # the single-pair call is the batch of one: same bits, and n_matched counts the keypoints that keep a depth
    k0, d0 = hk(kl)[3, :nl_h[3]], xl.cpu().numpy()[3, :nl_h[3]]
    k1, d1 = hk(kr)[3, :int(nr.cpu().numpy()[3])], xr.cpu().numpy()[3, :int(nr.cpu().numpy()[3])]
    u1, z1 = P.compute_stereo_matches(exl, exr, k0, d0, k1, d1, bf, b, frame=3)
    assert np.array_equal(u1.view(np.uint32), ur_h[3, :len(k0)].view(np.uint32)) and np.array_equal(z1.view(np.uint32), z_h[3, :len(k0)].view(np.uint32))
    # guards: keypoints that did not come from this extractor (octave out of range, window outside the level) are left unmatched
    bad = k0.copy()
    bad["octave"][::3] = 9; bad["octave"][1::3] = -1
    bad["x"][2::3] = 2.0; bad["y"][2::3] = 1.0
    u2, z2 = P.compute_stereo_matches(exl, exr, bad, d0, k1, d1, bf, b, frame=3)
    assert np.all(u2[::3] == -1) and np.all(u2[1::3] == -1) and np.all(z2[2::3] == -1)


def test_feature_vector_indices_are_checked(P, ctx, vocs):
    """feat[] entries index the descriptor / keypoint arrays on the device: anything outside [0, n) is refused on the host."""
    _, vo, vg = vocs
    d1, k1, d2, k2 = make_pair(31, 300, 280, 0.08, 0.0, KP=P.KP_DTYPE)
    _, sg = sides(P, vo, vg, d1, k1, d2, k2, 2, 31)
    m = P.ORBmatcher(ctx, 0.75, True)
    good = m.SearchByBoW(sg[0], sg[1])
    for side, bad_value in ((0, 300), (1, -1), (1, 1 << 30)):
        keep = sg[side].feat.copy()
        sg[side].feat[len(keep) // 2] = bad_value
        with pytest.raises(P.OrbxError):
            m.SearchByBoW(sg[0], sg[1])
        sg[side].feat[:] = keep
    again = m.SearchByBoW(sg[0], sg[1])
    assert np.array_equal(good[0], again[0]) and good[1] == again[1]


def test_search_by_bow_batch_equals_single_calls(P, ctx, vocs):
    """orbx_match_bow_batch: several keyframe pairs of different sizes / modes in one set of launches."""
    _, vo, vg = vocs
    cases = [(21, 900, 800, 0.08, 0.0, 1), (22, 1500, 1700, 0.1, 0.5, 2), (23, 30, 25, 0.05, 0.0, 2), (24, 2000, 2000, 0.1, 0.0, 4),
             (25, 600, 1200, 0.12, 0.8, 3)]
    for kf_kf in (False, True):
        pairs_g, want = [], []
        for seed, n1, n2, flip, dup, levelsup in cases:
            d1, k1, d2, k2 = make_pair(seed, n1, n2, flip, dup, KP=P.KP_DTYPE)
            so, sg = sides(P, vo, vg, d1, k1, d2, k2, levelsup, seed)
            pairs_g.append((sg[0], sg[1]))
            want.append(O.search_by_bow(so[0], so[1], 0.75, True, kf_kf))
        got = P.ORBmatcher(ctx, 0.75, True).SearchByBoWBatch(pairs_g, kf_kf)
        assert len(got) == len(want)
        for (wm, wn), (gm, gn) in zip(want, got):
            assert np.array_equal(wm, gm) and wn == gn
