"""CPU: the oracle (oracle/orb_extract.c, orb_match.c) against the UNMODIFIED reference sources compiled as
oracle/_ref/liborb_ref.so (`make -C oracle ref`; recipe and cv:: shim under oracle/).  This is what pins the oracle's
control flow -- cell loop, quadtree, ordering, every matcher loop, stereo, the DBoW2 descent -- to the reference itself.
The image primitives under the shim are the oracle's own, pinned to cv2 4.13 by tests/test_oracle_golden.py.

Quadtree tie-break: the reference orders equal-size nodes by heap address (src/ORBextractor.cc:935).  The default here is
the monotonic allocator model (address order == creation order), under which reference == oracle bit for bit;
test_real_heap_* measures what the process's real heap does instead."""
import ctypes
import os
import numpy as np
import pytest
import synth, oracle_lib as O, ref_lib as R, matcher_scene

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built and /root/reference absent")

CASES = [("tum", 0, 640, 480, 1000, "texture"), ("euroc", 1, 752, 480, 1200, "texture"), ("kitti", 4, 1241, 376, 2000, "texture"),
         ("grad", 2, 320, 240, 500, "gradient"), ("noise", 3, 640, 480, 1000, "noise"), ("black", 0, 160, 120, 300, "black"),
         ("small", 6, 232, 224, 200, "texture"), ("blocks", 8, 640, 480, 1000, "blocks"), ("hd", 9, 1920, 1080, 4000, "texture"),
         ("odd", 10, 333, 251, 700, "texture")]


# "small": the smallest shape every level of which still has a FAST cell.  Below it (a level narrower than 62 px) the
# reference divides by zero at src/ORBextractor.cc:1080 and dies in vector::resize; the oracle and the GPU path define
# such a level as "no keypoints" (tests/golden/extract_small.npz is 96x80), which therefore has no reference counterpart.


def _same_kps(a, b):
    return len(a) == len(b) and a.tobytes() == b.tobytes()


def test_ctor_tables_equal_reference():
    L = R.lib()
    for nf, sc, nl, ini, mn in [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (1200, 1.2, 8, 20, 7), (4000, 1.2, 8, 20, 7), (500, 1.5, 5, 12, 5), (300, 1.1, 12, 30, 10)]:
        f = [np.zeros(16, np.float32) for _ in range(4)]
        nfeat, umax = np.zeros(16, np.int32), np.zeros(16, np.int32)
        L.ref_extractor_tables(nf, sc, nl, ini, mn, *[a.ctypes.data for a in f], nfeat.ctypes.data, umax.ctypes.data)
        p = O.params(nf, sc, nl, ini, mn)
        for got, want in zip(f, (p.scale, p.inv_scale, p.sigma2, p.inv_sigma2)):
            assert np.array_equal(got[:nl].view(np.uint32), np.array(want[:nl], np.float32).view(np.uint32))
        assert list(nfeat[:nl]) == list(p.nfeat_level[:nl]) and list(umax) == list(p.umax)


@pytest.mark.parametrize("name,seed,w,h,nf,kind", CASES)
def test_extract_equals_reference(name, seed, w, h, nf, kind):
    """ORBextractor::operator() of the reference == orc_extract: keypoints as raw bytes, descriptors, every pyramid level
    (and the 19-px reflect-101 frame of mvImagePyramid against orc_border_reflect101)."""
    img = synth.frame(seed, w, h, kind)
    p = O.params(nf)
    ok, od, opyr = O.extract(p, img, want_pyr=True)
    R.set_allocator(1)
    rk, rd, rpyr, rb = R.extract(img, nf, want_pyr=True, want_bordered=True)
    assert _same_kps(ok, rk), (len(ok), len(rk))
    assert np.array_equal(od, rd)
    assert np.array_equal(opyr, rpyr)
    po = pb = 0
    for lw, lh in O.level_sizes(p, w, h):
        want = O.border(opyr[po:po + lw * lh].reshape(lh, lw), 19); po += lw * lh
        got = rb[pb:pb + (lw + 38) * (lh + 38)].reshape(lh + 38, lw + 38); pb += (lw + 38) * (lh + 38)
        assert np.array_equal(got, want)
    if kind != "black":
        assert len(ok) > 0


def test_extract_strided_input_and_other_parameters():
    img = synth.frame(31, 700, 500)
    view = img[7:487, 13:653]                                  # a 640x480 window with a 700-byte row stride
    R.set_allocator(1)
    for nf, sc, nl, ini, mn in [(1000, 1.2, 8, 20, 7), (800, 1.5, 5, 12, 5), (600, 1.1, 10, 25, 9)]:
        p = O.params(nf, sc, nl, ini, mn)
        ok, od = O.extract(p, np.ascontiguousarray(view))
        cap = nf + 64 * nl + 1024
        kps, desc = np.empty(cap, O.KP_DTYPE), np.empty((cap, 32), np.uint8)
        n = R.lib().ref_extract(nf, sc, nl, ini, mn, view.ctypes.data, 640, 480, view.strides[0], kps.ctypes.data, desc.ctypes.data, cap, None, None)
        assert n == len(ok) and kps[:n].tobytes() == ok.tobytes() and np.array_equal(desc[:n], od)


def test_distribute_octree_equals_reference():
    """DistributeOctTree on raw candidate lists (incl. more candidates than wanted, fewer, wide and tall windows)."""
    R.set_allocator(1)
    rng = np.random.default_rng(5)
    for n, W, H, N in [(5000, 608, 448, 217), (300, 608, 448, 500), (2000, 1209, 344, 434), (1500, 500, 700, 150), (1, 100, 100, 10), (40, 64, 48, 40)]:
        k = np.zeros(n, O.KP_DTYPE)
        k["x"] = rng.integers(0, W, n).astype(np.float32); k["y"] = rng.integers(0, H, n).astype(np.float32)
        k["response"] = rng.integers(7, 120, n).astype(np.float32); k["size"] = 7; k["angle"] = -1; k["class_id"] = -1
        want = O.distribute(k, 16, 16 + W, 16, 16 + H, N)
        got = R.distribute(k, 16, 16 + W, 16, 16 + H, N)
        assert _same_kps(want, got), (n, W, H, N, len(want), len(got))


@pytest.mark.parametrize("w,h,nf,seed", [(752, 480, 1200, 3), (1241, 376, 2000, 5), (640, 480, 1000, 17), (501, 333, 800, 23)])
def test_stereo_frame_equals_reference(w, h, nf, seed):
    """Frame::Frame(imLeft, imRight, ...) of the reference (two extractions, ComputeStereoMatches, src/Frame.cc:119-217,
    :831-1128) == orc_extract x2 + orc_stereo_matches: mvuRight and mvDepth as raw bits."""
    left, right = synth.stereo_pair(seed, w, h)
    p = O.params(nf)
    okl, odl, pyl = O.extract(p, left, want_pyr=True)
    okr, odr, pyr = O.extract(p, right, want_pyr=True)
    bf, fx = 386.1448, 718.856
    want_u, want_z, nv = O.stereo_matches(okl, odl, okr, odr, pyl, pyr, O.level_sizes(p, w, h), np.array(p.scale[:8], np.float32),
                                          np.array(p.inv_scale[:8], np.float32), bf, float(np.float32(bf) / np.float32(fx)))
    cap = nf + 2048
    kL, kR, kU = (np.empty(cap, O.KP_DTYPE) for _ in range(3))
    dL, dR = np.empty((cap, 32), np.uint8), np.empty((cap, 32), np.uint8)
    uR, z = np.empty(cap, np.float32), np.empty(cap, np.float32)
    nL, nR = ctypes.c_int(), ctypes.c_int()
    R.set_allocator(1)
    n = R.lib().ref_stereo_frame(nf, 1.2, 8, 20, 7, left.ctypes.data, right.ctypes.data, w, h, fx, 718.856, 607.1928, 185.2157, None, 0,
                                 bf, 35.0, kL.ctypes.data, dL.ctypes.data, ctypes.byref(nL), kR.ctypes.data, dR.ctypes.data,
                                 ctypes.byref(nR), cap, uR.ctypes.data, z.ctypes.data, kU.ctypes.data)
    assert n >= 0
    assert _same_kps(kL[:nL.value], okl) and _same_kps(kR[:nR.value], okr)
    assert np.array_equal(dL[:nL.value], odl) and np.array_equal(dR[:nR.value], odr)
    assert _same_kps(kU[:nL.value], okl)                                   # zero distortion: mvKeysUn == mvKeys
    assert np.array_equal(uR[:nL.value].view(np.uint32), want_u.view(np.uint32))
    assert np.array_equal(z[:nL.value].view(np.uint32), want_z.view(np.uint32))
    assert n == int((want_z > 0).sum()) and n > 100


@pytest.mark.parametrize("seed,n1,n2", [(11, 1500, 1300), (12, 700, 900), (13, 2000, 1900), (14, 300, 350)])
def test_matcher_scene_equals_reference(tmp_path, seed, n1, n2):
    """Every ORBmatcher method of the reference (src/ORBmatcher.cc, unmodified, over the reference's own Frame / KeyFrame /
    MapPoint / Map) on one scene file == the oracle's loops (+ a Python replay of Fuse's map updates)."""
    def run(i, o):
        assert R.lib().ref_run_matcher_scene(i.encode(), o.encode()) == 0
    R.set_allocator(0)
    matcher_scene.check_scene(run, tmp_path, seed, n1, n2)


def test_descriptor_distance_equals_reference():
    d = synth.descriptors(3, 400)
    for i in range(0, 400, 2):
        assert R.lib().ref_descriptor_distance(d[i].ctypes.data, d[i + 1].ctypes.data) == O.lib().orc_descriptor_distance(d[i].ctypes.data, d[i + 1].ctypes.data)


def test_real_heap_tie_break_is_measured():
    """With the process's real heap the reference may order equal-size quadtree nodes differently (address reuse), which
    changes a few keypoints per frame; everything that is not a tie must still agree.  tools/quadtree_ties.py turns this
    into the table of DESIGN.md section 2."""
    img = synth.frame(0, 640, 480)
    ok, od = O.extract(O.params(1000), img)
    R.set_allocator(0)
    rk, rd = R.extract(img, 1000)
    R.set_allocator(1)
    common = len(set(map(bytes, ok)) & set(map(bytes, rk)))
    assert abs(len(rk) - len(ok)) <= 24 and common >= 0.95 * len(ok), (len(ok), len(rk), common)
    both = {bytes(k): bytes(d) for k, d in zip(ok, od)}
    assert all(both[bytes(k)] == bytes(d) for k, d in zip(rk, rd) if bytes(k) in both)   # same keypoint -> same descriptor


def _ref_vocab(tmp_path, v, text):
    path = tmp_path / "voc.txt"
    # no trailing newline: the reference's `while(!f.eof()) getline` loop (TemplatedVocabulary.h:1425) would otherwise parse
    # one more, empty line with an uninitialised parent id (undefined behaviour; the product's loader skips empty lines)
    path.write_bytes((text if isinstance(text, bytes) else text.encode()).rstrip(b"\n"))
    h = R.lib().ref_vocab_load_text(str(path).encode())
    assert h
    os.unlink(path)
    return h


def _check_vocab(h, v, desc, levelsups):
    """DBoW2's own loadFromTextFile + transform (TemplatedVocabulary.h:1385-1487, :1145-1306) against orc_bow_transform,
    orc_feature_vector and orc_bow_vector on the flattened tree."""
    L = R.lib()
    vo = O.VocabArrays(v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    info = [ctypes.c_int() for _ in range(6)]
    L.ref_vocab_info(h, *[ctypes.byref(x) for x in info])
    assert [x.value for x in info] == [v["k"], v["L"], int(v["leaf"].sum()), len(v["desc"]), 0, 0]
    n = len(desc)
    ww = np.zeros(int(v["leaf"].sum()), np.float64)
    ww[v["word_id"][v["leaf"] > 0]] = v["weight"][v["leaf"] > 0].astype(np.float64)
    for levelsup in levelsups:
        word, wt, node = O.bow_transform(vo, desc, levelsup)
        rw, rwt, rn = np.empty(n, np.int32), np.empty(n, np.float64), np.empty(n, np.int32)
        L.ref_vocab_transform_each(h, desc.ctypes.data, n, levelsup, rw.ctypes.data, rwt.ctypes.data, rn.ctypes.data)
        assert np.array_equal(word, rw) and np.array_equal(node, rn) and np.array_equal(wt.astype(np.float64), rwt)
        nid, noff, feat = O.feature_vector(node, wt)
        ids, vals = O.bow_vector(word, ww)
        bi, bv = np.empty(n, np.int32), np.empty(n, np.float64)
        fn, fo, ff = np.empty(n + 1, np.int32), np.empty(n + 2, np.int32), np.empty(n + 1, np.int32)
        nn = ctypes.c_int()
        nb = L.ref_vocab_transform(h, desc.ctypes.data, n, levelsup, bi.ctypes.data, bv.ctypes.data, fn.ctypes.data, fo.ctypes.data,
                                   ff.ctypes.data, ctypes.byref(nn))
        assert nb == len(ids) and np.array_equal(bi[:nb], ids) and np.array_equal(bv[:nb].view(np.uint64), vals.view(np.uint64))
        assert nn.value == len(nid) and np.array_equal(fn[:nn.value], nid) and np.array_equal(fo[:nn.value + 1], noff)
        assert np.array_equal(ff[:noff[-1]], feat)
    # L1Scoring::score (ScoringObject.cpp:28-61) between two halves
    w1, _, _ = O.bow_transform(vo, desc[:n // 2], 4); w2, _, _ = O.bow_transform(vo, desc[n // 2:], 4)
    i1, v1 = O.bow_vector(w1, ww); i2, v2 = O.bow_vector(w2, ww)
    s = L.ref_vocab_score(h, i1.ctypes.data, v1.ctypes.data, len(i1), i2.ctypes.data, v2.ctypes.data, len(i2))
    assert np.float64(s).view(np.uint64) == np.float64(O.bow_score_l1(i1, v1, i2, v2)).view(np.uint64)


def _voc_queries(v, seed, n):
    """Descriptors near random leaves (so that descents are meaningful) plus pure noise."""
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(v["leaf"] > 0)
    near = synth.descriptors(seed, n * 3 // 4, 0.06, v["desc"][rng.choice(leaves, n * 3 // 4)])
    return np.ascontiguousarray(np.concatenate([near, synth.descriptors(seed + 1, n - len(near))]))


def test_vocabulary_small_equals_reference(tmp_path):
    v = synth.vocabulary(0, 10, 3)                          # depth-first node order, arbitrary float weights, stop words
    h = _ref_vocab(tmp_path, v, synth.vocabulary_text(v))
    _check_vocab(h, v, _voc_queries(v, 5, 3000), (0, 1, 2, 3, 4))
    R.lib().ref_vocab_free(h)
    v = synth.vocabulary_bfs(3, 9, 4)                       # breadth-first order, fixed-width text
    h = _ref_vocab(tmp_path, v, synth.vocabulary_text_fixed(v))
    _check_vocab(h, v, _voc_queries(v, 6, 2000), (0, 2, 4, 5))
    R.lib().ref_vocab_free(h)


def test_vocabulary_orbvoc_size_equals_reference(tmp_path):
    """configs[3]: k = 10, L = 6 (1 111 111 nodes, 10^6 words, a 170 MB text file) through DBoW2's own loader and descent:
    10 k descriptors at levelsup 4 (BoW mode, <= 100 nodes) and 6 (one node: brute force)."""
    v = synth.vocabulary_bfs(0, 10, 6)
    R.set_allocator(0)
    h = _ref_vocab(tmp_path, v, synth.vocabulary_text_fixed(v))
    _check_vocab(h, v, _voc_queries(v, 7, 10000), (4, 6, 0))
    R.lib().ref_vocab_free(h)


def test_extract_fuzz_sizes_and_parameters():
    """Random frame sizes (odd widths included), feature budgets, scale factors, level counts and FAST thresholds: the oracle
    equals the reference's ORBextractor::operator() byte for byte on each.  Sizes keep every level at least 62 px wide / high
    (below that the reference divides by zero at src/ORBextractor.cc:1080)."""
    rng = np.random.default_rng(2024)
    R.set_allocator(1)
    kinds = ["texture", "noise", "blocks", "gradient"]
    done = 0
    while done < 14:
        sc = float(rng.choice([1.1, 1.2, 1.25, 1.4, 1.5, 2.0]))
        nl = int(rng.integers(2, 9))
        w, h = int(rng.integers(150, 900)), int(rng.integers(150, 700))
        if min(w, h) / sc ** (nl - 1) < 70:
            continue
        nf = int(rng.integers(100, 2500))
        ini = int(rng.integers(8, 40)); mn = int(rng.integers(3, ini + 1))
        img = synth.frame(int(rng.integers(0, 10000)), w, h, kinds[done % 4])
        p = O.params(nf, sc, nl, ini, mn)
        ok, od = O.extract(p, img)
        rk, rd = R.extract(img, nf, sc, nl, ini, mn)
        assert _same_kps(ok, rk) and np.array_equal(od, rd), (w, h, nf, sc, nl, ini, mn, len(ok), len(rk))
        done += 1
