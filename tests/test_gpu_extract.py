"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle and the committed cv2 goldens.
Bit-exact on x, y, size, angle, response, octave and every descriptor byte."""
import os
import numpy as np
import pytest
import synth, oracle_lib as O
from gpu_util import pkg, assert_kps_equal

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def P():
    return pkg()


def test_ctor_tables_match_oracle(P):
    for nf, sc, nl in [(1000, 1.2, 8), (2000, 1.2, 8), (1200, 1.2, 8), (4000, 1.2, 8), (500, 1.5, 5), (300, 2.0, 4)]:
        ex = P.ORBextractor(nf, sc, nl, 20, 7, max_w=64, max_h=64)
        p = O.params(nf, sc, nl)
        assert ex.GetLevels() == nl and np.float32(sc) == np.float32(ex.GetScaleFactor())
        assert np.array_equal(ex.GetScaleFactors().view(np.uint32), np.array(p.scale[:nl], np.float32).view(np.uint32))
        assert np.array_equal(ex.GetInverseScaleFactors().view(np.uint32), np.array(p.inv_scale[:nl], np.float32).view(np.uint32))
        assert np.array_equal(ex.GetScaleSigmaSquares().view(np.uint32), np.array(p.sigma2[:nl], np.float32).view(np.uint32))
        assert np.array_equal(ex.GetInverseScaleSigmaSquares().view(np.uint32), np.array(p.inv_sigma2[:nl], np.float32).view(np.uint32))
        assert list(ex.features_per_level()) == list(p.nfeat_level[:nl])
        ex.close()


def test_pyramid_levels_bit_exact(P):
    img = synth.frame(7, 640, 480)
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=640, max_h=480)
    ex(img)
    p = O.params(1000)
    _, _, pyr = O.extract(p, img, want_pyr=True)
    o = 0
    for l, (w, h) in enumerate(O.level_sizes(p, 640, 480)):
        want = pyr[o:o + w * h].reshape(h, w)
        o += w * h
        assert np.array_equal(ex.pyramid_level(l), want), l
        assert np.array_equal(ex.mvImagePyramid[l], want)
        assert np.array_equal(ex.pyramid_level(l, border=19), O.border(want, 19)), ("border", l)


def test_bordered_pyramid_in_one_pass_and_matcher_context(P):
    """orbx_pyramid_bordered (what the C++ class mirrors into mvImagePyramid) == the per-level call, for a frame in the middle
    of a batch and an unaligned width; and orbx_create_matcher gives a context on which matcher calls work and extraction
    calls are refused."""
    import ctypes
    for w, h, nf in ((640, 480, 1000), (333, 251, 700)):
        imgs = np.stack([synth.frame(40 + i, w, h) for i in range(3)])
        ex = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=3)
        ex.extract_batch(imgs)
        for frame in (0, 2):
            got = ex.pyramid_bordered(frame, 19)
            for l in range(8):
                assert np.array_equal(got[l], ex.pyramid_level(l, frame=frame, border=19)), (w, frame, l)
        p = O.params(nf)
        _, _, pyr = O.extract(p, imgs[2], want_pyr=True)
        lw, lh = O.level_sizes(p, w, h)[0]
        assert np.array_equal(ex.pyramid_bordered(2, 19)[0], O.border(pyr[:lw * lh].reshape(lh, lw), 19))
    L = P.lib()
    hm = ctypes.c_void_p()
    assert L.orbx_create_matcher(ctypes.byref(hm), 0) == 0

    class M:                                            # just enough of the extractor wrapper for hamming_best2
        _h, _lib = hm, L

        def _check(self, rc, what):
            assert rc == 0, (what, rc)
    q, t = synth.descriptors(1, 300), synth.descriptors(2, 500)
    want = O.hamming_best2(q, t)
    got = P.hamming_best2(M(), q, t)
    assert all(np.array_equal(a, b) for a, b in zip(want, got))
    kps = np.empty(2000, P.KP_DTYPE); desc = np.empty((2000, 32), np.uint8); n = ctypes.c_int()
    img = synth.frame(1, 64, 64)
    rc = L.orbx_extract(hm, img.ctypes.data_as(ctypes.c_void_p), 64, 64, 64, 0, kps.ctypes.data_as(ctypes.c_void_p),
                        desc.ctypes.data_as(ctypes.c_void_p), 2000, ctypes.byref(n))
    assert rc == -4                                     # ORBX_E_CAPACITY: no extraction arenas on a matcher context
    L.orbx_destroy(hm)


CASES = ["tum", "euroc", "kitti", "grad", "noise", "black", "small", "blocks"]


@pytest.mark.parametrize("name", CASES)
def test_extract_matches_golden_and_oracle(P, name):
    g = np.load(os.path.join(G, f"extract_{name}.npz"))
    w, h, nf = int(g["w"]), int(g["h"]), int(g["nfeatures"])
    img = synth.frame(int(g["seed"]), w, h, str(g["kind"]))
    ex = P.ORBextractor(nf, 1.2, 8, 20, 7, max_w=w, max_h=h)
    kps, desc = ex(img)
    okps, odesc = O.extract(O.params(nf), img)
    assert_kps_equal(kps, okps, desc, odesc, name)
    kp = g["kp"]
    assert len(kps) == len(kp)
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(kps[f].view(np.uint32), kp[:, i].view(np.uint32)), f
    assert np.array_equal(desc, g["desc"])


def test_empty_image_leaves_outputs(P):
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)
    kps, desc = ex(np.zeros((0, 0), np.uint8))
    assert len(kps) == 0 and desc.shape == (0, 32)


def test_strided_input_and_size_change(P):
    ex = P.ORBextractor(800, 1.2, 8, 20, 7, max_w=800, max_h=600)
    big = synth.frame(9, 800, 600)
    for (x0, y0, w, h) in [(0, 0, 800, 600), (13, 7, 500, 333), (100, 50, 641, 479)]:
        view = big[y0:y0 + h, x0:x0 + w]          # non-contiguous rows
        kps, desc = ex(view)
        okps, odesc = O.extract(O.params(800), np.ascontiguousarray(view))
        assert_kps_equal(kps, okps, desc, odesc, (w, h))


def test_batch_equals_single(P):
    B = 6
    imgs = np.stack([synth.frame(100 + i, 640, 480, "texture" if i % 3 else "noise") for i in range(B)])
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=640, max_h=480, max_batch=B)
    res = ex.extract_batch(imgs)
    p = O.params(1000)
    for b in range(B):
        okps, odesc = O.extract(p, imgs[b])
        assert_kps_equal(res[b][0], okps, res[b][1], odesc, b)


def test_throughput_launch_shapes_equal_latency_shapes(P):
    """A batch that fills the GPU takes other launch shapes than a single frame (16-row resize threads, 128-thread FAST and
    quadtree CTAs, blur in line); both must give the oracle's result."""
    distinct = [synth.frame(300 + i, 640, 480, "texture" if i % 4 else "noise") for i in range(8)]
    imgs = np.stack([distinct[i % 8] for i in range(96)])
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=640, max_h=480, max_batch=96)
    res = ex.extract_batch(imgs)
    p = O.params(1000)
    want = [O.extract(p, im) for im in distinct]
    for b in range(96):
        assert_kps_equal(res[b][0], want[b % 8][0], res[b][1], want[b % 8][1], b)
    one = ex.extract_batch(imgs[:1])
    assert_kps_equal(one[0][0], want[0][0], one[0][1], want[0][1], "single")


def test_two_callers_two_handles(P):
    """The end-to-end pattern of bench.py: two host threads, one extractor handle each, batches in flight at the same time
    (the handles share the device, nothing else); every result equals the single-caller result."""
    import threading
    imgs = [np.stack([synth.frame(100 * t + s, 320, 240) for s in range(6)]) for t in range(2)]
    exs = [P.ORBextractor(500, 1.2, 8, 20, 7, max_w=320, max_h=240, max_batch=6) for _ in range(2)]
    want = [exs[0].extract_batch(im) for im in imgs]
    got = [[None] * 4 for _ in range(2)]

    def work(t):
        for r in range(4):
            got[t][r] = exs[t].extract_batch(imgs[t])
    th = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    for t in range(2):
        for r in range(4):
            for (k0, d0), (k1, d1) in zip(want[t], got[t][r]):
                assert_kps_equal(k1, k0, d1, d0, (t, r))


def test_configs_geometry_1080p(P):
    img = synth.frame(5, 1920, 1080)
    ex = P.ORBextractor(4000, 1.2, 8, 20, 7, max_w=1920, max_h=1080)
    kps, desc = ex(img)
    okps, odesc = O.extract(O.params(4000), img)
    assert_kps_equal(kps, okps, desc, odesc, "1080p")


def test_1080p_batch_of_64_at_full_size(P):
    """BASELINE.json configs[4] at its full size: 64 frames of 1920x1080 / 4000 features in one batch (512-thread quadtree CTAs,
    the staged resize on every level), every frame against the oracle (8 distinct frames, tiled)."""
    distinct = [synth.frame(900 + i, 1920, 1080, "texture" if i % 4 else "noise") for i in range(8)]
    imgs = np.stack([distinct[i % 8] for i in range(64)])
    ex = P.ORBextractor(4000, 1.2, 8, 20, 7, max_w=1920, max_h=1080, max_batch=64)
    res = ex.extract_batch(imgs)
    p = O.params(4000)
    want = [O.extract(p, im) for im in distinct]
    for b in range(64):
        assert_kps_equal(res[b][0], want[b % 8][0], res[b][1], want[b % 8][1], b)
    ex.close()


def test_idempotent_and_deterministic(P):
    img = synth.frame(3, 752, 480)
    ex = P.ORBextractor(1200, 1.2, 8, 20, 7, max_w=752, max_h=480)
    a = ex(img)
    for _ in range(3):
        b = ex(img)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_device_sincosf_and_atan2_exact(P):
    """Device glibc-sincosf port vs the oracle restatement on 64 Mi consecutive floats per band, and
    fastAtan2 vs the oracle on random integer moments."""
    import ctypes
    L = P.lib()
    n = 1 << 24
    lib = O.lib()
    for first in (0x00000000, 0x3A000000, 0x3F000000, 0x3F490000, 0x40000000, 0x40400000, 0x40C00000 - n + 0x100000):
        s = np.empty(n, np.float32); c = np.empty(n, np.float32)
        assert L.orbx_test_sincosf(first, n, s.ctypes.data, c.ctypes.data) == 0
        x = (np.arange(n, dtype=np.uint32) + np.uint32(first)).view(np.float32)
        idx = np.random.default_rng(first).integers(0, n, 20000)
        for i in idx[:4000]:
            assert np.float32(lib.orc_sinf(float(x[i]))).view(np.uint32) == s[i].view(np.uint32), (first, i)
            assert np.float32(lib.orc_cosf(float(x[i]))).view(np.uint32) == c[i].view(np.uint32), (first, i)
    rng = np.random.default_rng(1)
    yx = rng.integers(-3000000, 3000000, size=(200000, 2)).astype(np.float32)
    yx[:4] = [[0, 0], [0, 5], [5, 0], [-5, -5]]
    out = np.empty(len(yx), np.float32)
    assert L.orbx_test_fast_atan2(yx.ctypes.data, len(yx), out.ctypes.data) == 0
    want = np.array([lib.orc_fast_atan2(float(a), float(b)) for a, b in yx[:50000]], np.float32)
    assert np.array_equal(out[:50000].view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("pad_x,pad_w", [(1, 3), (4, 8), (16, 32), (0, 0)])
def test_device_input_any_alignment(P, pad_x, pad_w):
    """Device-resident frames whose base pointer / row pitch are not 16- or even 4-byte aligned take the plain-load
    staging paths of every kernel (no bulk copies, generic resize); results must not change."""
    import torch
    B, w, h = 3, 640, 480
    imgs = np.stack([synth.frame(40 + i, w, h) for i in range(B)])
    big = torch.zeros((B, h + 2, w + pad_w), dtype=torch.uint8, device="cuda:0")
    view = big[:, 1:h + 1, pad_x:pad_x + w]
    view.copy_(torch.from_numpy(imgs).to("cuda:0"))
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=w, max_h=h, max_batch=B)
    cap = ex.max_keypoints()
    kps = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda:0")
    desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda:0")
    n = torch.zeros((B,), dtype=torch.int32, device="cuda:0")
    ex.extract_batch_device(view, kps, desc, n)
    ex.sync()
    p = O.params(1000)
    for b in range(B):
        nb = int(n[b])
        k = np.frombuffer(kps[b, :nb].cpu().numpy().tobytes(), dtype=P.KP_DTYPE)
        okps, odesc = O.extract(p, imgs[b])
        assert_kps_equal(k, okps, desc[b, :nb].cpu().numpy(), odesc, (pad_x, pad_w, b))


def test_cvt_gray_matches_golden_and_oracle(P):
    """cvtColor -> gray in front of the extractor (src/Tracking.cc:252-278): cv2 golden, oracle on odd sizes / strides."""
    import os
    ex = P.ORBextractor(100, 1.2, 8, 20, 7, max_w=64, max_h=64)
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cvt_gray.npz"))
    rng = np.random.default_rng(77)
    inputs = {"rgb": rng.integers(0, 256, (37, 61, 3), dtype=np.uint8), "bgr": rng.integers(0, 256, (40, 64, 3), dtype=np.uint8),
              "rgba": rng.integers(0, 256, (33, 50, 4), dtype=np.uint8), "bgra": rng.integers(0, 256, (16, 31, 4), dtype=np.uint8)}
    fm = {"rgb": P.FMT_RGB, "bgr": P.FMT_BGR, "rgba": P.FMT_RGBA, "bgra": P.FMT_BGRA}
    for k, im in inputs.items():
        assert np.array_equal(P.cvt_gray(ex, im, fm[k]), g[k]), k
    for (h, w, ch, pad) in ((480, 640, 3, 0), (376, 1241, 3, 5), (1, 1, 3, 0), (7, 3, 4, 2), (1080, 1920, 4, 0), (33, 1023, 3, 1)):
        buf = rng.integers(0, 256, (h, w + pad, ch), dtype=np.uint8)
        im = buf[:, :w]                                   # rows with a stride the kernel has to honour
        for bgr in (False, True):
            want = O.cvt_gray(np.ascontiguousarray(im), bgr)
            got = P.cvt_gray(ex, im, (2 if ch == 4 else 0) | int(bgr))
            assert np.array_equal(want, got), (h, w, ch, pad, bgr)


def test_undistort_keypoints_matches_golden_and_oracle(P):
    """Frame::UndistortKeyPoints (src/Frame.cc:725-781): cv2 golden + oracle on real extractor output."""
    import os
    from test_oracle_golden import UNDISTORT, undistort_points
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=640, max_h=480)
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "undistort.npz"))
    k = undistort_points()
    for name, (K, D) in UNDISTORT.items():
        out = P.undistort_keypoints(ex, k, K, D)
        assert np.array_equal(out["x"].view(np.uint32), g[name][:, 0].view(np.uint32)), name
        assert np.array_equal(out["y"].view(np.uint32), g[name][:, 1].view(np.uint32)), name
        assert np.array_equal(out["octave"], k["octave"]) and np.array_equal(out["angle"], k["angle"])
    kps, _ = ex(synth.frame(0, 640, 480))
    K, D = UNDISTORT["tum1"]
    assert np.array_equal(P.undistort_keypoints(ex, kps, K, D), O.undistort_keypoints(kps, K, D))
    assert np.array_equal(P.undistort_keypoints(ex, kps, K, (0.0, 0.1, 0.0, 0.0)), kps)
