"""ctypes binding of oracle/liborb_oracle.so (TEST INFRASTRUCTURE: the checker, never the product)."""
import ctypes
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
MAXL = 16


class Keypoint(ctypes.Structure):
    _fields_ = [("x", ctypes.c_float), ("y", ctypes.c_float), ("size", ctypes.c_float), ("angle", ctypes.c_float),
                ("response", ctypes.c_float), ("octave", ctypes.c_int32), ("class_id", ctypes.c_int32)]


KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


class Params(ctypes.Structure):
    _fields_ = [("nfeatures", ctypes.c_int), ("nlevels", ctypes.c_int), ("ini_th", ctypes.c_int), ("min_th", ctypes.c_int),
                ("scale_factor", ctypes.c_float),
                ("scale", ctypes.c_float * MAXL), ("inv_scale", ctypes.c_float * MAXL),
                ("sigma2", ctypes.c_float * MAXL), ("inv_sigma2", ctypes.c_float * MAXL),
                ("nfeat_level", ctypes.c_int * MAXL), ("umax", ctypes.c_int * 16), ("blur_ksize", ctypes.c_int)]


_lib = None


def build():
    subprocess.run(["make", "-C", ODIR, "-s"], check=True)


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ODIR, "liborb_oracle.so")
        srcs = [os.path.join(ODIR, f) for f in os.listdir(ODIR) if f.endswith((".c", ".h", ".inc"))]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            build()
        L = ctypes.CDLL(so)
        u8p = ctypes.c_void_p
        L.orc_params_init.argtypes = [ctypes.POINTER(Params), ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_level_size.argtypes = [ctypes.POINTER(Params), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
        L.orc_resize_linear.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t]
        L.orc_border_reflect101.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_size_t, ctypes.c_int]
        L.orc_gauss_blur.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_size_t, ctypes.c_int]
        L.orc_fast_atan2.restype = ctypes.c_float
        L.orc_fast_atan2.argtypes = [ctypes.c_float, ctypes.c_float]
        L.orc_sinf.restype = ctypes.c_float
        L.orc_sinf.argtypes = [ctypes.c_float]
        L.orc_cosf.restype = ctypes.c_float
        L.orc_cosf.argtypes = [ctypes.c_float]
        L.orc_fast_roi.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_int, u8p, ctypes.c_int]
        L.orc_detect_level.argtypes = [ctypes.POINTER(Params), u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_int]
        L.orc_distribute_octree.argtypes = [u8p, ctypes.c_int] + [ctypes.c_int] * 5 + [u8p, ctypes.c_int]
        L.orc_ic_angle.restype = ctypes.c_float
        L.orc_ic_angle.argtypes = [u8p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.orc_brief.argtypes = [u8p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_float, u8p]
        L.orc_extract.argtypes = [ctypes.POINTER(Params), u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, u8p, ctypes.c_int, u8p]
        L.orc_descriptor_distance.argtypes = [u8p, u8p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def params(nfeatures=1000, scale=1.2, nlevels=8, ini=20, mn=7):
    p = Params()
    rc = lib().orc_params_init(ctypes.byref(p), nfeatures, scale, nlevels, ini, mn)
    assert rc == 0
    return p


def level_sizes(p, w, h):
    out = []
    for l in range(p.nlevels):
        a, b = ctypes.c_int(), ctypes.c_int()
        lib().orc_level_size(ctypes.byref(p), l, w, h, ctypes.byref(a), ctypes.byref(b))
        out.append((a.value, b.value))
    return out


def resize(src, dw, dh):
    src = np.ascontiguousarray(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border(src, b=19):
    src = np.ascontiguousarray(src)
    h, w = src.shape
    dst = np.empty((h + 2 * b, w + 2 * b), np.uint8)
    lib().orc_border_reflect101(_p(src), w, h, src.strides[0], _p(dst), dst.strides[0], b)
    return dst


def blur(src, ksize=9):
    src = np.ascontiguousarray(src)
    dst = np.empty_like(src)
    lib().orc_gauss_blur(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0], ksize)
    return dst


def cvt_gray(img, bgr=False):
    """cv::cvtColor(..2GRAY) restatement; img: (h, w, 3|4) uint8."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w, ch = img.shape
    out = np.empty((h, w), np.uint8)
    L = lib()
    L.orc_cvt_gray.restype = None
    L.orc_cvt_gray.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t]
    L.orc_cvt_gray(_p(img), w, h, w * ch, ch, int(bgr), _p(out), w)
    return out


def undistort_keypoints(kps, K, dist):
    """Frame::UndistortKeyPoints; K = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3])."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    d = np.ascontiguousarray(dist, np.float32)
    out = np.empty(max(len(kps), 1), KP_DTYPE)
    L = lib()
    L.orc_undistort_keypoints.restype = None
    L.orc_undistort_keypoints.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_float] * 4 + [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    L.orc_undistort_keypoints(_p(kps), len(kps), *[float(np.float32(x)) for x in K], _p(d), len(d), _p(out))
    return out[:len(kps)]


def fast_roi(roi, th):
    roi = np.ascontiguousarray(roi)
    cap = roi.size
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast_roi(_p(roi), roi.shape[1], roi.shape[0], roi.strides[0], th, _p(out), cap)
    return out[:n].copy()


def detect_level(p, img):
    img = np.ascontiguousarray(img)
    cap = img.size
    out = np.empty(cap, KP_DTYPE)
    n = lib().orc_detect_level(ctypes.byref(p), _p(img), img.shape[1], img.shape[0], img.strides[0], _p(out), cap)
    return out[:n].copy()


def distribute(kps, minx, maxx, miny, maxy, N):
    kps = np.ascontiguousarray(kps)
    out = np.empty(max(len(kps), 1), KP_DTYPE)
    n = lib().orc_distribute_octree(_p(kps), len(kps), minx, maxx, miny, maxy, N, _p(out), len(out))
    assert n >= 0
    return out[:n].copy()


def extract(p, img, cap=None, want_pyr=False):
    img = np.ascontiguousarray(img)
    h, w = img.shape
    cap = cap or (p.nfeatures + 64 * p.nlevels + 1024)
    kps = np.empty(cap, KP_DTYPE)
    desc = np.empty((cap, 32), np.uint8)
    pyr = None
    if want_pyr:
        pyr = np.empty(sum(a * b for a, b in level_sizes(p, w, h)), np.uint8)
    n = lib().orc_extract(ctypes.byref(p), _p(img), w, h, img.strides[0], _p(kps), _p(desc), cap, _p(pyr) if want_pyr else None)
    assert 0 <= n <= cap, n
    if want_pyr:
        return kps[:n].copy(), desc[:n].copy(), pyr
    return kps[:n].copy(), desc[:n].copy()


# ---------------------------------------------------------------- matcher side (oracle/orb_match.c)
class Vocab(ctypes.Structure):
    _fields_ = [("k", ctypes.c_int), ("L", ctypes.c_int), ("nnodes", ctypes.c_int), ("child_off", ctypes.c_void_p),
                ("child", ctypes.c_void_p), ("desc", ctypes.c_void_p), ("word_id", ctypes.c_void_p), ("weight", ctypes.c_void_p)]


class BowFrame(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int), ("desc", ctypes.c_void_p), ("keys", ctypes.c_void_p), ("valid", ctypes.c_void_p),
                ("uright", ctypes.c_void_p), ("nnodes", ctypes.c_int), ("node_id", ctypes.c_void_p),
                ("node_off", ctypes.c_void_p), ("feat", ctypes.c_void_p)]


class Grid(ctypes.Structure):
    _fields_ = [("min_x", ctypes.c_float), ("min_y", ctypes.c_float), ("inv_w", ctypes.c_float), ("inv_h", ctypes.c_float)]


PROJ_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("ur", "<f4"), ("view_cos", "<f4"), ("angle", "<f4"), ("level", "<i4"),
                       ("desc", "u1", (32,))])
assert PROJ_DTYPE.itemsize == 56
GRID_COLS, GRID_ROWS = 64, 48


def _mlib():
    L = lib()
    if getattr(L, "_match_ready", False):
        return L
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    L.orc_three_maxima.argtypes = [vp, ci, vp, vp, vp]
    L.orc_bow_transform.argtypes = [ctypes.POINTER(Vocab), vp, ci, ci, vp, vp, vp]
    L.orc_feature_vector.argtypes = [vp, vp, ci, vp, vp, vp]
    L.orc_search_by_bow_kf_f.argtypes = [ctypes.POINTER(BowFrame), ctypes.POINTER(BowFrame), cf, ci, vp]
    L.orc_search_by_bow_kf_kf.argtypes = [ctypes.POINTER(BowFrame), ctypes.POINTER(BowFrame), cf, ci, vp]
    L.orc_search_for_triangulation.argtypes = [ctypes.POINTER(BowFrame), ctypes.POINTER(BowFrame), vp, cf, cf, vp, vp, ci, ci, vp]
    L.orc_assign_grid.argtypes = [vp, ci, ctypes.POINTER(Grid), vp, vp]
    L.orc_features_in_area.argtypes = [vp, ctypes.POINTER(Grid), vp, vp, cf, cf, cf, ci, ci, vp]
    L.orc_search_by_projection_points.argtypes = [vp, ci, vp, vp, ci, vp, vp, ctypes.POINTER(Grid), vp, vp, vp, cf, cf, vp, vp]
    L.orc_search_by_projection_frame.argtypes = [vp, ci, vp, vp, ci, vp, vp, ctypes.POINTER(Grid), vp, vp, vp, cf, ci, ci, vp, vp]
    L.orc_search_by_projection_reloc.argtypes = [vp, ci, vp, vp, ci, vp, ctypes.POINTER(Grid), vp, vp, vp, cf, ci, ci, vp]
    L.orc_search_by_projection_sim3.argtypes = [vp, ci, vp, vp, ci, vp, ctypes.POINTER(Grid), vp, vp, vp, ci, vp]
    L.orc_search_for_initialization.argtypes = [vp, vp, ci, vp, vp, vp, ci, ctypes.POINTER(Grid), vp, vp, ci, cf, ci, vp]
    L.orc_window_best.argtypes = [vp, ci, vp, vp, vp, ci, vp, ctypes.POINTER(Grid), vp, vp, cf, vp, ci, vp, vp]
    L.orc_window_best.restype = None
    L.orc_distinctive_descriptors.argtypes = [vp, vp, ci, vp]
    L.orc_distinctive_descriptors.restype = None
    L.orc_bow_vector.argtypes = [vp, ci, vp, ci, vp, vp]
    L.orc_bow_score_l1.argtypes = [vp, vp, ci, vp, vp, ci]
    L.orc_bow_score_l1.restype = ctypes.c_double
    L.orc_stereo_matches.argtypes = [vp, vp, ci, vp, vp, ci, vp, vp, vp, vp, vp, vp, vp, cf, cf, vp, vp]
    L.orc_hamming_best2.argtypes = [vp, vp, vp, vp, vp, vp, ci, vp, vp, vp, vp]
    L._match_ready = True
    return L


def hamming_best2(q, t, q_seg=None, q_idx=None, t_seg=None, t_idx=None, t_valid=None):
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
    t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    if q_seg is None:
        q_seg, t_seg = np.array([0, len(q)], np.int32), np.array([0, len(t)], np.int32)
    q_seg, t_seg = np.ascontiguousarray(q_seg, np.int32), np.ascontiguousarray(t_seg, np.int32)
    opt = lambda a, dt: None if a is None else np.ascontiguousarray(a, dt)
    q_idx, t_idx, t_valid = opt(q_idx, np.int32), opt(t_idx, np.int32), opt(t_valid, np.uint8)
    n = len(q) if q_idx is None else len(q_idx)
    bi, bd, sd = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.int32)
    pp = lambda a: None if a is None else _p(a)
    _mlib().orc_hamming_best2(_p(q), _p(t), _p(q_seg), pp(q_idx), _p(t_seg), pp(t_idx), len(q_seg) - 1, pp(t_valid), _p(bi), _p(bd), _p(sd))
    return bi, bd, sd


def three_maxima(counts):
    counts = np.ascontiguousarray(counts, np.int32)
    a, b, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _mlib().orc_three_maxima(_p(counts), len(counts), ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
    return a.value, b.value, c.value


class VocabArrays:
    """Flattened vocabulary tree (see orc_vocab)."""

    def __init__(self, k, L, child_off, child, desc, word_id, weight):
        self.k, self.L = k, L
        self.child_off = np.ascontiguousarray(child_off, np.int32)
        self.child = np.ascontiguousarray(child, np.int32)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.word_id = np.ascontiguousarray(word_id, np.int32)
        self.weight = np.ascontiguousarray(weight, np.float32)
        self.nnodes = len(self.word_id)

    def c(self):
        return Vocab(self.k, self.L, self.nnodes, self.child_off.ctypes.data, self.child.ctypes.data, self.desc.ctypes.data,
                     self.word_id.ctypes.data, self.weight.ctypes.data)


def bow_transform(voc, desc, levelsup):
    desc = np.ascontiguousarray(desc, np.uint8)
    n = len(desc)
    word, weight, node = np.empty(n, np.int32), np.empty(n, np.float32), np.empty(n, np.int32)
    v = voc.c()
    _mlib().orc_bow_transform(ctypes.byref(v), _p(desc), n, levelsup, _p(word), _p(weight), _p(node))
    return word, weight, node


def feature_vector(node, weight):
    n = len(node)
    node = np.ascontiguousarray(node, np.int32)
    weight = np.ascontiguousarray(weight, np.float32)
    nid, noff, feat = np.empty(n + 1, np.int32), np.empty(n + 2, np.int32), np.empty(n + 1, np.int32)
    nn = _mlib().orc_feature_vector(_p(node), _p(weight), n, _p(nid), _p(noff), _p(feat))
    return nid[:nn].copy(), noff[:nn + 1].copy(), feat[:noff[nn]].copy()


class Side:
    """One (key)frame for the BoW matchers; keeps the numpy arrays alive."""

    def __init__(self, desc, keys, valid, uright, node_id, node_off, feat):
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.keys = np.ascontiguousarray(keys)
        self.valid = np.ascontiguousarray(valid, np.uint8)
        self.uright = None if uright is None else np.ascontiguousarray(uright, np.float32)
        self.node_id = np.ascontiguousarray(node_id, np.int32)
        self.node_off = np.ascontiguousarray(node_off, np.int32)
        self.feat = np.ascontiguousarray(feat, np.int32)

    def c(self):
        return BowFrame(len(self.desc), self.desc.ctypes.data, self.keys.ctypes.data, self.valid.ctypes.data,
                        None if self.uright is None else self.uright.ctypes.data, len(self.node_id),
                        self.node_id.ctypes.data, self.node_off.ctypes.data, self.feat.ctypes.data)


def search_by_bow(s1, s2, nnratio, check_ori, kf_kf):
    a, b = s1.c(), s2.c()
    out = np.empty(len(s1.desc) if kf_kf else len(s2.desc), np.int32)
    fn = _mlib().orc_search_by_bow_kf_kf if kf_kf else _mlib().orc_search_by_bow_kf_f
    n = fn(ctypes.byref(a), ctypes.byref(b), nnratio, int(check_ori), _p(out))
    return out, n


def search_for_triangulation(s1, s2, F12, ex, ey, scale2, sigma2, only_stereo, check_ori):
    a, b = s1.c(), s2.c()
    F12 = np.ascontiguousarray(F12, np.float32).reshape(9)
    scale2 = np.ascontiguousarray(scale2, np.float32)
    sigma2 = np.ascontiguousarray(sigma2, np.float32)
    out = np.empty(len(s1.desc), np.int32)
    n = _mlib().orc_search_for_triangulation(ctypes.byref(a), ctypes.byref(b), _p(F12), ex, ey, _p(scale2), _p(sigma2),
                                             int(only_stereo), int(check_ori), _p(out))
    return out, n


def assign_grid(keys, grid):
    keys = np.ascontiguousarray(keys)
    off = np.empty(GRID_COLS * GRID_ROWS + 1, np.int32)
    idx = np.empty(max(len(keys), 1), np.int32)
    g = Grid(*grid)
    _mlib().orc_assign_grid(_p(keys), len(keys), ctypes.byref(g), _p(off), _p(idx))
    return off, idx[:off[-1]].copy()


def search_by_projection(pts, keys, desc, uright, scale, grid, occupied, th, nnratio=None, motion=0, check_ori=True, transient=None):
    """nnratio given -> SearchByProjection(Frame, MapPoints) (a15); else the Frame/Frame variant (a16).
    transient[i] != 0: point i has Observations() == 0, its match does not hide the keypoint from later points."""
    tr = None if transient is None else np.ascontiguousarray(transient, np.uint8)
    pts = np.ascontiguousarray(pts)
    keys = np.ascontiguousarray(keys)
    desc = np.ascontiguousarray(desc, np.uint8)
    scale = np.ascontiguousarray(scale, np.float32)
    off, idx = assign_grid(keys, grid)
    idx = np.ascontiguousarray(np.concatenate([idx, np.zeros(1, np.int32)]))
    g = Grid(*grid)
    out = np.empty(len(keys), np.int32)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    occ = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    if nnratio is not None:
        n = _mlib().orc_search_by_projection_points(_p(pts), len(pts), _p(keys), _p(desc), len(keys), None if ur is None else _p(ur),
                                                    _p(scale), ctypes.byref(g), _p(off), _p(idx), None if occ is None else _p(occ),
                                                    th, nnratio, _p(out), None if tr is None else _p(tr))
    else:
        n = _mlib().orc_search_by_projection_frame(_p(pts), len(pts), _p(keys), _p(desc), len(keys), None if ur is None else _p(ur),
                                                   _p(scale), ctypes.byref(g), _p(off), _p(idx), None if occ is None else _p(occ),
                                                   th, motion, int(check_ori), _p(out), None if tr is None else _p(tr))
    return out, n


def _proj_common(pts, keys, desc, scale, grid, occupied):
    pts = np.ascontiguousarray(pts)
    keys = np.ascontiguousarray(keys)
    desc = np.ascontiguousarray(desc, np.uint8)
    scale = np.ascontiguousarray(scale, np.float32)
    off, idx = assign_grid(keys, grid)
    idx = np.ascontiguousarray(np.concatenate([idx, np.zeros(1, np.int32)]))
    occ = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    return pts, keys, desc, scale, off, idx, Grid(*grid), occ, np.empty(len(keys), np.int32)


def search_by_projection_reloc(pts, keys, desc, scale, grid, occupied, th, orb_dist, check_ori=True):
    """SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, ORBdist) (a17), after the caller's projection."""
    pts, keys, desc, scale, off, idx, g, occ, out = _proj_common(pts, keys, desc, scale, grid, occupied)
    n = _mlib().orc_search_by_projection_reloc(_p(pts), len(pts), _p(keys), _p(desc), len(keys), _p(scale), ctypes.byref(g), _p(off),
                                               _p(idx), None if occ is None else _p(occ), th, int(orb_dist), int(check_ori), _p(out))
    return out, n


def search_by_projection_sim3(pts, keys, desc, scale, grid, occupied, th):
    """SearchByProjection(KeyFrame, Scw, vpPoints, vpMatched, th) (a18), after the caller's Sim3 projection."""
    pts, keys, desc, scale, off, idx, g, occ, out = _proj_common(pts, keys, desc, scale, grid, occupied)
    n = _mlib().orc_search_by_projection_sim3(_p(pts), len(pts), _p(keys), _p(desc), len(keys), _p(scale), ctypes.byref(g), _p(off),
                                              _p(idx), None if occ is None else _p(occ), int(th), _p(out))
    return out, n


def window_best(pts, keys, desc, uright, scale, grid, th, inv_sigma2, th_dist):
    """Candidate loops of Fuse x2 / SearchBySim3 (independent queries).  Returns (best_idx, best_dist)."""
    pts, keys, desc, scale, off, idx, g, _, _ = _proj_common(pts, keys, desc, scale, grid, None)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    is2 = None if inv_sigma2 is None else np.ascontiguousarray(inv_sigma2, np.float32)
    bi, bd = np.empty(max(len(pts), 1), np.int32), np.empty(max(len(pts), 1), np.int32)
    _mlib().orc_window_best(_p(pts), len(pts), _p(keys), _p(desc), None if ur is None else _p(ur), len(keys), _p(scale), ctypes.byref(g),
                            _p(off), _p(idx), th, None if is2 is None else _p(is2), int(th_dist), _p(bi), _p(bd))
    return bi[:len(pts)], bd[:len(pts)]


def bow_vector(word, word_weight):
    """BowVector (ids ascending, L1-normalised double values) of the features' word ids; word_weight[id] = WordValue."""
    word = np.ascontiguousarray(word, np.int32); ww = np.ascontiguousarray(word_weight, np.float64)
    ids = np.empty(max(len(word), 1), np.int32); vals = np.empty(max(len(word), 1), np.float64)
    n = _mlib().orc_bow_vector(_p(word), len(word), _p(ww), len(ww), _p(ids), _p(vals))
    return ids[:n].copy(), vals[:n].copy()


def bow_score_l1(ids1, v1, ids2, v2):
    ids1 = np.ascontiguousarray(ids1, np.int32); ids2 = np.ascontiguousarray(ids2, np.int32)
    v1 = np.ascontiguousarray(v1, np.float64); v2 = np.ascontiguousarray(v2, np.float64)
    return _mlib().orc_bow_score_l1(_p(ids1), _p(v1), len(ids1), _p(ids2), _p(v2), len(ids2))


def distinctive_descriptors(desc, seg):
    """MapPoint::ComputeDistinctiveDescriptors for many map points: desc rows [seg[p], seg[p+1]) belong to point p."""
    desc = np.ascontiguousarray(desc, np.uint8); seg = np.ascontiguousarray(seg, np.int32)
    out = np.empty(max(len(seg) - 1, 1), np.int32)
    _mlib().orc_distinctive_descriptors(_p(desc), _p(seg), len(seg) - 1, _p(out))
    return out[:len(seg) - 1]


def search_for_initialization(keys1, desc1, prev_xy, keys2, desc2, grid, window, nnratio, check_ori=True):
    """SearchForInitialization (src/ORBmatcher.cc:580-748).  Returns (matches12, nmatches, updated prev_xy)."""
    keys1 = np.ascontiguousarray(keys1); keys2 = np.ascontiguousarray(keys2)
    desc1 = np.ascontiguousarray(desc1, np.uint8); desc2 = np.ascontiguousarray(desc2, np.uint8)
    prev = np.ascontiguousarray(prev_xy, np.float32).copy()
    off, idx = assign_grid(keys2, grid)
    idx = np.ascontiguousarray(np.concatenate([idx, np.zeros(1, np.int32)]))
    g = Grid(*grid)
    out = np.empty(max(len(keys1), 1), np.int32)
    n = _mlib().orc_search_for_initialization(_p(keys1), _p(desc1), len(keys1), _p(prev), _p(keys2), _p(desc2), len(keys2),
                                              ctypes.byref(g), _p(off), _p(idx), int(window), nnratio, int(check_ori), _p(out))
    return out[:len(keys1)], n, prev


def stereo_matches(kL, dL, kR, dR, pyrL, pyrR, sizes, scale, inv_scale, mbf, mb):
    kL, kR = np.ascontiguousarray(kL), np.ascontiguousarray(kR)
    dL, dR = np.ascontiguousarray(dL, np.uint8), np.ascontiguousarray(dR, np.uint8)
    w = np.array([s[0] for s in sizes], np.int32)
    h = np.array([s[1] for s in sizes], np.int32)
    off = np.zeros(len(sizes), np.uint64)
    o = 0
    for i, (a, b) in enumerate(sizes):
        off[i] = o
        o += a * b
    scale = np.ascontiguousarray(scale, np.float32)
    inv_scale = np.ascontiguousarray(inv_scale, np.float32)
    ur, z = np.empty(len(kL), np.float32), np.empty(len(kL), np.float32)
    n = _mlib().orc_stereo_matches(_p(kL), _p(dL), len(kL), _p(kR), _p(dR), len(kR), _p(pyrL), _p(pyrR), _p(w), _p(h), _p(off),
                                   _p(scale), _p(inv_scale), mbf, mb, _p(ur), _p(z))
    return ur, z, n
