"""CPU: the oracle (oracle/*.c) against golden vectors made with the real cv2 4.13 / libm
primitives (tools/gen_golden.py).  This is what pins the oracle; see oracle/orb_oracle.h."""
import hashlib, os
import numpy as np
import pytest
import synth, oracle_lib as O

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def prim():
    return np.load(os.path.join(G, "primitives.npz"))


def test_resize_matches_cv2(prim):
    img = synth.frame(11, 213, 160, "blocks")
    for key in prim.files:
        if key.startswith("resize_"):
            dw, dh = map(int, key[7:].split("x"))
            assert np.array_equal(O.resize(img, dw, dh), prim[key]), key


def test_blur_border_match_cv2(prim):
    img = synth.frame(11, 213, 160, "blocks")
    assert np.array_equal(O.blur(img, 9), prim["blur9"])
    assert np.array_equal(O.blur(img, 7), prim["blur7"])
    assert np.array_equal(O.border(img, 19), prim["border19"])


def test_fast_atan2_matches_cv2(prim):
    out = np.array([O.lib().orc_fast_atan2(float(y), float(x)) for y, x in prim["atan2_in"]], np.float32)
    assert np.array_equal(out.view(np.uint32), prim["atan2_out"].view(np.uint32))


def test_fast_roi_matches_cv2(prim):
    for th in (7, 20):
        for seed, kind in [(21, "texture"), (22, "noise"), (23, "gradient")]:
            roi = synth.frame(seed, 43, 41, kind)
            assert np.array_equal(O.fast_roi(roi, th), prim[f"fast_{kind}_{th}"]), (kind, th)


CASES = ["tum", "euroc", "kitti", "grad", "noise", "black", "small", "blocks"]


@pytest.mark.parametrize("name", CASES)
def test_extract_matches_cv2_pipeline(name):
    g = np.load(os.path.join(G, f"extract_{name}.npz"))
    img = synth.frame(int(g["seed"]), int(g["w"]), int(g["h"]), str(g["kind"]))
    assert sha(img) == str(g["image_sha"]), "synthetic generator drifted"
    p = O.params(int(g["nfeatures"]))
    kps, desc, pyr = O.extract(p, img, want_pyr=True)
    o = 0
    for (w, h), want in zip(O.level_sizes(p, img.shape[1], img.shape[0]), g["pyr_sha"]):
        assert sha(pyr[o:o + w * h]) == str(want)
        o += w * h
    kp = g["kp"]
    assert len(kps) == len(kp)
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(kps[f].view(np.uint32), kp[:, i].view(np.uint32)), f
    assert np.array_equal(kps["octave"], g["octave"])
    assert np.array_equal(desc, g["desc"])
    assert (kps["class_id"] == -1).all()


def test_ctor_tables():
    p = O.params(1000)
    assert list(p.nfeat_level)[:8] == [217, 181, 151, 126, 105, 87, 73, 60]
    assert list(p.umax) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert O.level_sizes(p, 640, 480) == [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]
    assert list(O.params(4000).nfeat_level)[:8] == [869, 724, 603, 503, 419, 349, 291, 242]


def test_cvt_gray_matches_cv2():
    """cvtColor RGB/BGR/RGBA/BGRA -> GRAY (src/Tracking.cc:252-278): the oracle against cv2 4.13 outputs (tests/golden/cvt_gray.npz)."""
    g = np.load(os.path.join(G, "cvt_gray.npz"))
    rng = np.random.default_rng(77)
    inputs = {"rgb": rng.integers(0, 256, (37, 61, 3), dtype=np.uint8), "bgr": rng.integers(0, 256, (40, 64, 3), dtype=np.uint8),
              "rgba": rng.integers(0, 256, (33, 50, 4), dtype=np.uint8), "bgra": rng.integers(0, 256, (16, 31, 4), dtype=np.uint8)}
    for k, im in inputs.items():
        assert np.array_equal(O.cvt_gray(im, bgr=k.startswith("bgr")), g[k]), k
    r_, g_, b_ = np.meshgrid(np.arange(256), np.arange(256), np.arange(0, 256, 3), indexing="ij")
    sweep = np.stack([r_, g_, b_], -1).astype(np.uint8).reshape(256, -1, 3)
    assert sha(O.cvt_gray(sweep)) == str(g["sweep_sha"])


UNDISTORT = {"tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
             "tum2": ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
             "four": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05))}


def undistort_points():
    rng = np.random.default_rng(78)
    pts = np.stack([rng.random(3000) * 640, rng.random(3000) * 480], 1).astype(np.float32)
    pts[:4] = [[0, 0], [639, 479], [318.64304, 255.313989], [-20, 500]]
    k = np.zeros(len(pts), O.KP_DTYPE)
    k["x"], k["y"], k["octave"], k["angle"] = pts[:, 0], pts[:, 1], np.arange(len(pts)) % 8, 12.5
    return k


def test_undistort_keypoints_matches_cv2():
    """Frame::UndistortKeyPoints (src/Frame.cc:725-781): the oracle against cv2.undistortPoints outputs (tests/golden/undistort.npz)."""
    g = np.load(os.path.join(G, "undistort.npz"))
    k = undistort_points()
    for name, (K, D) in UNDISTORT.items():
        out = O.undistort_keypoints(k, K, D)
        assert np.array_equal(out["x"].view(np.uint32), g[name][:, 0].view(np.uint32)), name
        assert np.array_equal(out["y"].view(np.uint32), g[name][:, 1].view(np.uint32)), name
        assert np.array_equal(out["octave"], k["octave"]) and np.array_equal(out["angle"], k["angle"])
    same = O.undistort_keypoints(k, UNDISTORT["tum1"][0], (0.0, 0.1, 0, 0))
    assert np.array_equal(same, k)
