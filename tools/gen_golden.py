#!/usr/bin/env python3
"""Generate tests/golden/*.npz with the REAL third-party primitives (cv2 4.13, host libm).

Run in the build container only (cv2 present):  python tools/gen_golden.py
Fixtures pin the oracle (oracle/) to executable copies of the arithmetic the reference
calls: cv::resize, cv::GaussianBlur, cv::FAST per ROI, cv::fastAtan2, cv::copyMakeBorder,
and the whole operator() through tools/pyref_orb.py (cv2 primitives + restated control flow).
Inputs come from tests/synth.py seeds, so only outputs are stored.
"""
import os, sys, hashlib
import numpy as np, cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth, pyref_orb as R
G = os.path.join(ROOT, "tests", "golden")
os.makedirs(G, exist_ok=True)

def sha(a): return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()

# --- cvtColor -> gray (separate small fixture; `gen_golden.py --only-cvt` regenerates just this one) ---
def cvt_inputs():
    rng = np.random.default_rng(77)
    return {"rgb": rng.integers(0, 256, (37, 61, 3), dtype=np.uint8), "bgr": rng.integers(0, 256, (40, 64, 3), dtype=np.uint8),
            "rgba": rng.integers(0, 256, (33, 50, 4), dtype=np.uint8), "bgra": rng.integers(0, 256, (16, 31, 4), dtype=np.uint8)}
codes = {"rgb": cv2.COLOR_RGB2GRAY, "bgr": cv2.COLOR_BGR2GRAY, "rgba": cv2.COLOR_RGBA2GRAY, "bgra": cv2.COLOR_BGRA2GRAY}
cvt = {k: cv2.cvtColor(v, codes[k]) for k, v in cvt_inputs().items()}
r_, g_, b_ = np.meshgrid(np.arange(256), np.arange(256), np.arange(0, 256, 3), indexing="ij")
sweep = np.stack([r_, g_, b_], -1).astype(np.uint8).reshape(256, -1, 3)      # every (R, G) with every third B
cvt["sweep_sha"] = np.array(sha(cv2.cvtColor(sweep, cv2.COLOR_RGB2GRAY)))
np.savez_compressed(os.path.join(G, "cvt_gray.npz"), **cvt)
# --- Frame::UndistortKeyPoints = cv2.undistortPoints(pts, K, D, None, K) (outputs only; inputs are seeded) ---
def undistort_cases():
    rng = np.random.default_rng(78)
    pts = np.stack([rng.random(3000) * 640, rng.random(3000) * 480], 1).astype(np.float32)
    pts[:4] = [[0, 0], [639, 479], [318.64304, 255.313989], [-20, 500]]
    return pts, {"tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
                 "tum2": ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
                 "four": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05))}
und = {}
upts, ucases = undistort_cases()
for name, (Kp, Dp) in ucases.items():
    Km = np.array([[Kp[0], 0, Kp[2]], [0, Kp[1], Kp[3]], [0, 0, 1]], np.float32)
    Dm = np.array(Dp, np.float32).reshape(-1, 1)
    und[name] = cv2.undistortPoints(upts.reshape(-1, 1, 2).copy(), Km, Dm, None, Km).reshape(-1, 2)
np.savez_compressed(os.path.join(G, "undistort.npz"), **und)
if "--only-cvt" in sys.argv or "--only-small" in sys.argv:
    sys.exit(0)

# --- primitives ---
prim = {}
img = synth.frame(11, 213, 160, "blocks")
for (dw, dh) in [(178, 133), (200, 150), (107, 80), (300, 231)]:
    prim[f"resize_{dw}x{dh}"] = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
prim["blur9"] = cv2.GaussianBlur(img, (9, 9), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
prim["blur7"] = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
prim["border19"] = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
rng = np.random.default_rng(5)
yx = rng.integers(-2000000, 2000000, size=(4000, 2)).astype(np.float32)
yx[:8] = [[0, 0], [0, 1], [1, 0], [0, -1], [-1, 0], [5, 5], [-5, 5], [3, -3]]
prim["atan2_in"] = yx
prim["atan2_out"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in yx], np.float32)
for th in (7, 20):
    for k, (seed, kind) in enumerate([(21, "texture"), (22, "noise"), (23, "gradient")]):
        roi = synth.frame(seed, 43, 41, kind)
        kp = cv2.FastFeatureDetector_create(th, True).detect(roi)
        prim[f"fast_{kind}_{th}"] = np.array([[int(q.pt[0]), int(q.pt[1]), int(q.response)] for q in kp], np.int32).reshape(-1, 3)
np.savez_compressed(os.path.join(G, "primitives.npz"), **prim)

# --- whole operator() ---
pat = R.load_pattern(os.path.join(ROOT, "oracle", "brief_pattern.inc"))
cases = [("tum", 0, 640, 480, 1000, "texture"), ("euroc", 1, 752, 480, 1200, "texture"),
         ("kitti", 4, 1241, 376, 2000, "texture"), ("grad", 2, 320, 240, 500, "gradient"),
         ("noise", 3, 640, 480, 1000, "noise"), ("black", 0, 160, 120, 300, "black"),
         ("small", 6, 96, 80, 200, "texture"), ("blocks", 8, 640, 480, 1000, "blocks")]
for name, seed, w, h, nf, kind in cases:
    im = synth.frame(seed, w, h, kind)
    kp, octv, desc, pyr, cands = R.extract(im, R.Params(nf, 1.2, 8, 20, 7), pat)
    np.savez_compressed(os.path.join(G, f"extract_{name}.npz"), seed=seed, w=w, h=h, nfeatures=nf, kind=kind,
                        image_sha=sha(im), kp=kp, octave=octv, desc=desc,
                        pyr_sha=np.array([sha(x) for x in pyr]), ncand=np.array([len(c) for c in cands]))
    print(name, len(kp), [len(c) for c in cands])
