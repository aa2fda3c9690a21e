"""Host-side mirror of ORB_SLAM2::ORBmatcher (include/ORBmatcher.h:48-264), the DBoW2 vocabulary transform and
Frame::ComputeStereoMatches over the C ABI (include/orbx.h).  SLAM objects are replaced by plain arrays:

  FrameView   what the matchers read from a Frame / KeyFrame: descriptors (N x 32), undistorted keypoints
              (KP_DTYPE), per-feature "has a good MapPoint" flags, mvuRight, and the FeatureVector in CSR form.

All compute runs in liborbx.so (CUDA); nothing here falls back to the CPU.
"""
import ctypes
import numpy as np
from .binding import KP_DTYPE, OrbxError, lib, _np

MAX_LEVELS = 16
BOW_KF_FRAME, BOW_KF_KF, BOW_TRIANGULATE = 0, 1, 2
PROJ_RATIO_SAME_LEVEL, PROJ_CHECK_URIGHT = 1, 2
FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48  # include/Frame.h:55,60

PROJ_QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("ur", "<f4"), ("angle", "<f4"),
                             ("min_level", "<i4"), ("max_level", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])
assert PROJ_QUERY_DTYPE.itemsize == 64


class _BowSide(ctypes.Structure):
    _fields_ = [("desc", ctypes.c_void_p), ("keys", ctypes.c_void_p), ("has_point", ctypes.c_void_p), ("uright", ctypes.c_void_p),
                ("n", ctypes.c_int), ("nnodes", ctypes.c_int), ("node_id", ctypes.c_void_p), ("node_off", ctypes.c_void_p),
                ("feat", ctypes.c_void_p)]


class _Triangulation(ctypes.Structure):
    _fields_ = [("F12", ctypes.c_float * 9), ("ex", ctypes.c_float), ("ey", ctypes.c_float),
                ("scale2", ctypes.c_float * MAX_LEVELS), ("sigma2_2", ctypes.c_float * MAX_LEVELS), ("only_stereo", ctypes.c_int)]


class _Grid(ctypes.Structure):
    _fields_ = [("min_x", ctypes.c_float), ("min_y", ctypes.c_float), ("inv_w", ctypes.c_float), ("inv_h", ctypes.c_float),
                ("cols", ctypes.c_int), ("rows", ctypes.c_int)]


_ready = False


def _mlib():
    global _ready
    L = lib()
    if _ready:
        return L
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    ip = ctypes.POINTER(ci)
    L.orbx_vocab_create.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, vp, ctypes.POINTER(vp)]
    L.orbx_vocab_load_text.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(vp)]
    L.orbx_vocab_destroy.argtypes = [vp]
    L.orbx_vocab_destroy.restype = None
    L.orbx_vocab_save_text.argtypes = [vp, ctypes.c_char_p]
    L.orbx_vocab_info.argtypes = [vp, ip, ip, ip, ip]
    L.orbx_vocab_modes.argtypes = [vp, ip, ip]
    L.orbx_bow_transform.argtypes = [vp, vp, vp, ci, ci, ci, vp, vp, vp]
    L.orbx_feature_vector.argtypes = [vp, vp, ci, vp, vp, vp, ip]
    L.orbx_match_bow.argtypes = [vp, ci, ctypes.POINTER(_BowSide), ctypes.POINTER(_BowSide), ctypes.POINTER(_Triangulation), cf, ci,
                                 ci, vp, ip]
    L.orbx_match_bow_batch.argtypes = [vp, ci, ci, ctypes.POINTER(ctypes.POINTER(_BowSide)), ctypes.POINTER(ctypes.POINTER(_BowSide)), vp, cf, ci, ci, vp, vp]
    L.orbx_assign_grid.argtypes = [vp, vp, ci, ctypes.POINTER(_Grid), ci, vp, vp]
    L.orbx_search_by_projection.argtypes = [vp, vp, ci, vp, vp, vp, ci, ctypes.POINTER(_Grid), vp, ci, ci, cf, ci, ci, vp, ip]
    L.orbx_search_for_initialization.argtypes = [vp, vp, vp, ci, vp, vp, vp, ci, ctypes.POINTER(_Grid), ci, cf, ci, ci, vp, ip]
    L.orbx_window_best.argtypes = [vp, vp, ci, vp, vp, vp, ci, ctypes.POINTER(_Grid), vp, ci, ci, vp, vp]
    L.orbx_distinctive_descriptors.argtypes = [vp, vp, vp, ci, ci, vp]
    L.orbx_vocab_set_word_weights.argtypes = [vp, vp, ci]
    L.orbx_bow_vector.argtypes = [vp, vp, vp, ci, ci, vp, vp, ip]
    L.orbx_bow_score_l1.argtypes = [vp, vp, vp, ci, vp, vp, vp, ci, ci, vp]
    L.orbx_stereo_match.argtypes = [vp, vp, ci, vp, vp, ci, vp, vp, ci, cf, cf, ci, vp, vp, ip]
    L.orbx_stereo_match_batch.argtypes = [vp, vp, ci, ci, vp, vp, vp, vp, vp, vp, ci, cf, cf, ci, vp, vp, vp]
    _ready = True
    return L


def _chk(ex, rc, what):
    if rc:
        L = lib()
        raise OrbxError(f"{what}: {L.orbx_strerror(rc).decode()} ({rc}) {L.orbx_last_cuda_error(ex._h).decode()}")


class Vocabulary:
    """ORBVocabulary (DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>): load + transform only."""

    def __init__(self, ex, k=None, L=None, child_off=None, child=None, desc=None, word_id=None, weight=None, path=None):
        self._ex, self._v = ex, ctypes.c_void_p()
        M = _mlib()
        if path is not None:
            _chk(ex, M.orbx_vocab_load_text(ex._h, path.encode(), ctypes.byref(self._v)), "orbx_vocab_load_text")
        else:
            a = [np.ascontiguousarray(child_off, np.int32), np.ascontiguousarray(child, np.int32), np.ascontiguousarray(desc, np.uint8),
                 np.ascontiguousarray(word_id, np.int32), np.ascontiguousarray(weight, np.float32)]
            _chk(ex, M.orbx_vocab_create(ex._h, k, L, len(a[3]), *[_np(x) for x in a], ctypes.byref(self._v)), "orbx_vocab_create")

    def info(self):
        v = [ctypes.c_int() for _ in range(4)]
        _mlib().orbx_vocab_info(self._v, *[ctypes.byref(x) for x in v])
        return tuple(x.value for x in v)  # k, L, nnodes, nwords

    def save_text(self, path):
        """TemplatedVocabulary::saveToTextFile (TemplatedVocabulary.h:1492-1517)."""
        _chk(self._ex, _mlib().orbx_vocab_save_text(self._v, path.encode()), "orbx_vocab_save_text")

    def modes(self):
        """(scoring, weighting) of the text header; (0, 0) = L1_NORM / TF_IDF."""
        a, b = ctypes.c_int(), ctypes.c_int()
        _mlib().orbx_vocab_modes(self._v, ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value

    def transform(self, desc, levelsup=4):
        """-> (word_id, weight, node_id) per feature; see feature_vector() for the FeatureVector."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        word, weight, node = np.empty(n, np.int32), np.empty(n, np.float32), np.empty(n, np.int32)
        _chk(self._ex, _mlib().orbx_bow_transform(self._ex._h, self._v, _np(desc), n, levelsup, 0, _np(word), _np(weight), _np(node)),
             "orbx_bow_transform")
        return word, weight, node

    def set_word_weights(self, weights):
        """Exact double weights per word id (WordValue); orbx_vocab_create takes floats, the text loader sets these itself."""
        w = np.ascontiguousarray(weights, np.float64)
        _chk(self._ex, _mlib().orbx_vocab_set_word_weights(self._v, _np(w), len(w)), "orbx_vocab_set_word_weights")

    def bow_vector(self, word):
        """BowVector of transform (TF-IDF weights, L1 norm) from the features' word ids -> (ids ascending, double values)."""
        word = np.ascontiguousarray(word, np.int32)
        ids, vals = np.empty(max(len(word), 1), np.int32), np.empty(max(len(word), 1), np.float64)
        n = ctypes.c_int()
        _chk(self._ex, _mlib().orbx_bow_vector(self._ex._h, self._v, _np(word), len(word), 0, _np(ids), _np(vals), ctypes.byref(n)),
             "orbx_bow_vector")
        return ids[:n.value].copy(), vals[:n.value].copy()

    def close(self):
        if self._v:
            _mlib().orbx_vocab_destroy(self._v)
            self._v = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def feature_vector(node, weight=None):
    node = np.ascontiguousarray(node, np.int32)
    n = len(node)
    w = None if weight is None else np.ascontiguousarray(weight, np.float32)
    nid, noff, feat = np.empty(n + 1, np.int32), np.empty(n + 2, np.int32), np.empty(n + 1, np.int32)
    nn = ctypes.c_int()
    rc = _mlib().orbx_feature_vector(_np(node), _np(w), n, _np(nid), _np(noff), _np(feat), ctypes.byref(nn))
    if rc:
        raise OrbxError(f"orbx_feature_vector: {rc}")
    return nid[:nn.value].copy(), noff[:nn.value + 1].copy(), feat[:noff[nn.value]].copy()


class FrameView:
    def __init__(self, desc, keys, has_point=None, uright=None, node_id=None, node_off=None, feat=None):
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        self.keys = np.ascontiguousarray(keys, KP_DTYPE)
        self.has_point = None if has_point is None else np.ascontiguousarray(has_point, np.uint8)
        self.uright = None if uright is None else np.ascontiguousarray(uright, np.float32)
        self.node_id = None if node_id is None else np.ascontiguousarray(node_id, np.int32)
        self.node_off = None if node_off is None else np.ascontiguousarray(node_off, np.int32)
        self.feat = None if feat is None else np.ascontiguousarray(feat, np.int32)

    @property
    def N(self):
        return len(self.desc)

    def _c(self):
        return _BowSide(self.desc.ctypes.data, self.keys.ctypes.data, None if self.has_point is None else self.has_point.ctypes.data,
                        None if self.uright is None else self.uright.ctypes.data, self.N, 0 if self.node_id is None else len(self.node_id),
                        None if self.node_id is None else self.node_id.ctypes.data, None if self.node_off is None else self.node_off.ctypes.data,
                        None if self.feat is None else self.feat.ctypes.data)


class ORBmatcher:
    """ORBmatcher(nnratio=0.6, checkOri=True), include/ORBmatcher.h:57."""
    TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30  # src/ORBmatcher.cc:56-58

    def __init__(self, ctx, nnratio=0.6, checkOri=True):
        self._ex, self.mfNNratio, self.mbCheckOrientation = ctx, float(nnratio), bool(checkOri)

    def _bow(self, mode, s1, s2, tri=None):
        a, b = s1._c(), s2._c()
        out = np.empty(s2.N if mode == BOW_KF_FRAME else s1.N, np.int32)
        n = ctypes.c_int()
        rc = _mlib().orbx_match_bow(self._ex._h, mode, ctypes.byref(a), ctypes.byref(b), None if tri is None else ctypes.byref(tri),
                                    self.mfNNratio, int(self.mbCheckOrientation), 0, _np(out), ctypes.byref(n))
        _chk(self._ex, rc, "orbx_match_bow")
        return out, n.value

    def SearchByBoWBatch(self, pairs, other_is_keyframe=False):
        """Many independent (KeyFrame, Frame/KeyFrame) pairs in one call (orbx_match_bow_batch): the keyframe-pair jobs of
        LocalMapping / LoopClosing.  pairs: list of (FrameView, FrameView).  Returns a list of (matches, n) identical to
        calling SearchByBoW on each pair."""
        mode = BOW_KF_KF if other_is_keyframe else BOW_KF_FRAME
        n = len(pairs)
        if n == 0:
            return []
        a = [p[0]._c() for p in pairs]
        b = [p[1]._c() for p in pairs]
        PS = ctypes.POINTER(_BowSide)
        pa = (PS * n)(*[ctypes.pointer(x) for x in a])
        pb = (PS * n)(*[ctypes.pointer(x) for x in b])
        outs = [np.empty(p[1].N if mode == BOW_KF_FRAME else p[0].N, np.int32) for p in pairs]
        pm = (ctypes.c_void_p * n)(*[o.ctypes.data for o in outs])
        nm = (ctypes.c_int * n)()
        rc = _mlib().orbx_match_bow_batch(self._ex._h, mode, n, pa, pb, None, self.mfNNratio, int(self.mbCheckOrientation), 0, pm, nm)
        _chk(self._ex, rc, "orbx_match_bow_batch")
        return [(outs[i], int(nm[i])) for i in range(n)]

    def SearchByBoW(self, kf, other, other_is_keyframe=False):
        """SearchByBoW(KeyFrame*, Frame&, ..) src/ORBmatcher.cc:255 -> (vpMapPointMatches as KF feature index per F feature, n);
        SearchByBoW(KeyFrame*, KeyFrame*, ..) :760 -> (vpMatches12 as KF2 feature index per KF1 feature, n)."""
        return self._bow(BOW_KF_KF if other_is_keyframe else BOW_KF_FRAME, kf, other)

    def SearchForTriangulation(self, kf1, kf2, F12, epipole, scale_factors2, level_sigma2_2, bOnlyStereo=False):
        """src/ORBmatcher.cc:915 -> (vMatches12, n); matched pairs are (i, vMatches12[i]) for vMatches12[i] >= 0."""
        t = _Triangulation()
        F = np.ascontiguousarray(F12, np.float32).reshape(9)
        for i in range(9):
            t.F12[i] = float(F[i])
        t.ex, t.ey = float(epipole[0]), float(epipole[1])
        for i, (s, g) in enumerate(zip(scale_factors2, level_sigma2_2)):
            t.scale2[i], t.sigma2_2[i] = float(s), float(g)
        t.only_stereo = int(bOnlyStereo)
        return self._bow(BOW_TRIANGULATE, kf1, kf2, t)

    def SearchByProjection(self, frame, queries, grid, th_dist, occupied=None, ratio_same_level=False, check_uright=True,
                           check_ori=None, grid_dims=(FRAME_GRID_COLS, FRAME_GRID_ROWS)):
        """The window search shared by the four SearchByProjection overloads after their projection step.
        queries: PROJ_QUERY_DTYPE array in the reference's point order; grid = (mnMinX, mnMinY, widthInv, heightInv).
        Returns (assign, n): assign[i] = query matched to frame keypoint i or -1."""
        q = np.ascontiguousarray(queries, PROJ_QUERY_DTYPE)
        g = _Grid(float(grid[0]), float(grid[1]), float(grid[2]), float(grid[3]), grid_dims[0], grid_dims[1])
        out = np.empty(frame.N, np.int32)
        n = ctypes.c_int()
        occ = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        mode = (PROJ_RATIO_SAME_LEVEL if ratio_same_level else 0) | (PROJ_CHECK_URIGHT if check_uright and frame.uright is not None else 0)
        co = self.mbCheckOrientation if check_ori is None else check_ori
        rc = _mlib().orbx_search_by_projection(self._ex._h, _np(q), len(q), _np(frame.keys), _np(frame.desc), _np(frame.uright), frame.N,
                                               ctypes.byref(g), _np(occ), mode, int(th_dist), self.mfNNratio, int(co), 0, _np(out),
                                               ctypes.byref(n))
        _chk(self._ex, rc, "orbx_search_by_projection")
        return out, n.value


    def SearchForInitialization(self, f1, f2, prev_matched, grid, window_size=100, grid_dims=(FRAME_GRID_COLS, FRAME_GRID_ROWS)):
        """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:580-748).  f1, f2: FrameView (mvKeysUn, mDescriptors);
        prev_matched: (N1, 2) float32 vbPrevMatched; grid: F2's (mnMinX, mnMinY, widthInv, heightInv).
        Returns (vnMatches12, nmatches, updated vbPrevMatched)."""
        prev = np.ascontiguousarray(prev_matched, np.float32).copy()
        assert prev.shape == (f1.N, 2)
        g = _Grid(float(grid[0]), float(grid[1]), float(grid[2]), float(grid[3]), grid_dims[0], grid_dims[1])
        out = np.empty(max(f1.N, 1), np.int32)
        n = ctypes.c_int()
        rc = _mlib().orbx_search_for_initialization(self._ex._h, _np(f1.keys), _np(f1.desc), f1.N, _np(prev), _np(f2.keys), _np(f2.desc),
                                                    f2.N, ctypes.byref(g), int(window_size), self.mfNNratio, int(self.mbCheckOrientation),
                                                    0, _np(out), ctypes.byref(n))
        _chk(self._ex, rc, "orbx_search_for_initialization")
        return out[:f1.N], n.value, prev


def window_best(ex, frame, queries, grid, th_dist, inv_sigma2=None, grid_dims=(FRAME_GRID_COLS, FRAME_GRID_ROWS)):
    """Candidate loops of ORBmatcher::Fuse x2 (src/ORBmatcher.cc:1139-1300 with inv_sigma2 = mvInvLevelSigma2, :1310-1458) and
    SearchBySim3 (:1472-1723): every query independently.  Returns (best_idx, best_dist)."""
    q = np.ascontiguousarray(queries, PROJ_QUERY_DTYPE)
    g = _Grid(float(grid[0]), float(grid[1]), float(grid[2]), float(grid[3]), grid_dims[0], grid_dims[1])
    bi, bd = np.empty(max(len(q), 1), np.int32), np.empty(max(len(q), 1), np.int32)
    is2 = None
    if inv_sigma2 is not None:
        is2 = np.zeros(16, np.float32); is2[:len(inv_sigma2)] = inv_sigma2
    rc = _mlib().orbx_window_best(ex._h, _np(q), len(q), _np(frame.keys), _np(frame.desc), _np(frame.uright), frame.N, ctypes.byref(g),
                                  _np(is2), int(th_dist), 0, _np(bi), _np(bd))
    _chk(ex, rc, "orbx_window_best")
    return bi[:len(q)], bd[:len(q)]


def bow_score_l1(ex, q_ids, q_vals, vectors):
    """L1Scoring::score (ScoringObject.cpp:28-61) of one BowVector against many: vectors = list of (ids, values)."""
    q_ids = np.ascontiguousarray(q_ids, np.int32); q_vals = np.ascontiguousarray(q_vals, np.float64)
    seg = np.concatenate([[0], np.cumsum([len(v[0]) for v in vectors])]).astype(np.int32)
    ids = np.ascontiguousarray(np.concatenate([np.asarray(v[0], np.int32) for v in vectors] + [np.zeros(0, np.int32)]), np.int32)
    vals = np.ascontiguousarray(np.concatenate([np.asarray(v[1], np.float64) for v in vectors] + [np.zeros(0, np.float64)]), np.float64)
    out = np.empty(max(len(vectors), 1), np.float64)
    _chk(ex, _mlib().orbx_bow_score_l1(ex._h, _np(q_ids), _np(q_vals), len(q_ids), _np(ids), _np(vals), _np(seg), len(vectors), 0, _np(out)),
         "orbx_bow_score_l1")
    return out[:len(vectors)]


def distinctive_descriptors(ex, desc, seg):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:366-448) for many map points per call."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    seg = np.ascontiguousarray(seg, np.int32)
    out = np.empty(max(len(seg) - 1, 1), np.int32)
    rc = _mlib().orbx_distinctive_descriptors(ex._h, _np(desc), _np(seg), len(seg) - 1, 0, _np(out))
    _chk(ex, rc, "orbx_distinctive_descriptors")
    return out[:len(seg) - 1]


def assign_features_to_grid(ex, keys, grid, grid_dims=(FRAME_GRID_COLS, FRAME_GRID_ROWS)):
    """Frame::AssignFeaturesToGrid (src/Frame.cc:398-422) -> CSR (cell_off, cell_idx), cell = ix*rows + iy."""
    keys = np.ascontiguousarray(keys, KP_DTYPE)
    g = _Grid(float(grid[0]), float(grid[1]), float(grid[2]), float(grid[3]), grid_dims[0], grid_dims[1])
    off = np.empty(grid_dims[0] * grid_dims[1] + 1, np.int32)
    idx = np.empty(max(len(keys), 1), np.int32)
    _chk(ex, _mlib().orbx_assign_grid(ex._h, _np(keys), len(keys), ctypes.byref(g), 0, _np(off), _np(idx)), "orbx_assign_grid")
    return off, idx[:off[-1]].copy()


def compute_stereo_matches(left, right, kps_l, desc_l, kps_r, desc_r, bf, b, frame=0):
    """Frame::ComputeStereoMatches (src/Frame.cc:831-1128): left/right are the two ORBextractor objects that just
    extracted kps_l / kps_r (their device pyramids are used).  Returns (mvuRight, mvDepth)."""
    kps_l, kps_r = np.ascontiguousarray(kps_l, KP_DTYPE), np.ascontiguousarray(kps_r, KP_DTYPE)
    desc_l, desc_r = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    ur, z = np.empty(len(kps_l), np.float32), np.empty(len(kps_l), np.float32)
    n = ctypes.c_int()
    rc = _mlib().orbx_stereo_match(left._h, right._h, frame, _np(kps_l), _np(desc_l), len(kps_l), _np(kps_r), _np(desc_r), len(kps_r),
                                   float(bf), float(b), 0, _np(ur), _np(z), ctypes.byref(n))
    _chk(left, rc, "orbx_stereo_match")
    return ur, z


def compute_stereo_matches_batch(left, right, kps_l, desc_l, n_l, kps_r, desc_r, n_r, bf, b, frame0=0, device=False):
    """Frame::ComputeStereoMatches for many stereo pairs in one call (orbx_stereo_match_batch).  kps / desc: (pairs, cap, ..)
    blocks and per-pair counts exactly as extract_batch wrote them for the left and the right batch -- numpy arrays, or torch
    CUDA tensors with device=True (the call then chains behind the two extractions without a host round trip).
    Returns (mvuRight, mvDepth) as (pairs, cap) arrays / tensors and the per-pair number of keypoints that got a depth."""
    pairs, cap = int(kps_l.shape[0]), int(kps_l.shape[1])
    nm = (ctypes.c_int32 * max(pairs, 1))()
    if device:
        import torch
        ur = torch.empty((pairs, cap), dtype=torch.float32, device=kps_l.device)
        z = torch.empty((pairs, cap), dtype=torch.float32, device=kps_l.device)
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        rc = _mlib().orbx_stereo_match_batch(left._h, right._h, frame0, pairs, P(kps_l), P(desc_l), P(n_l), P(kps_r), P(desc_r), P(n_r), cap,
                                             float(bf), float(b), 3, P(ur), P(z), nm)
    else:
        kps_l, kps_r = np.ascontiguousarray(kps_l, KP_DTYPE), np.ascontiguousarray(kps_r, KP_DTYPE)
        desc_l, desc_r = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
        n_l, n_r = np.ascontiguousarray(n_l, np.int32), np.ascontiguousarray(n_r, np.int32)
        ur, z = np.full((pairs, cap), -1, np.float32), np.full((pairs, cap), -1, np.float32)
        rc = _mlib().orbx_stereo_match_batch(left._h, right._h, frame0, pairs, _np(kps_l), _np(desc_l), _np(n_l), _np(kps_r), _np(desc_r),
                                             _np(n_r), cap, float(bf), float(b), 0, _np(ur), _np(z), nm)
    _chk(left, rc, "orbx_stereo_match_batch")
    return ur, z, np.array(nm[:pairs], np.int32)

