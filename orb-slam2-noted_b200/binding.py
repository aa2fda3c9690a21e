"""ctypes binding of liborbx.so + the reference-shaped Python host classes."""
import ctypes
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
IN_DEVICE, OUT_DEVICE = 1, 2
_lib = None


class OrbxError(RuntimeError):
    pass


def lib_path():
    """liborbx.so next to this file; env ORBX_LIB names another build of the same library (kernel A/B runs)."""
    return os.environ.get("ORBX_LIB") or os.path.join(HERE, "liborbx.so")


def lib():
    """Load liborbx.so; fail loudly when it has not been built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise OrbxError(f"{p} is missing: run `python __graft_entry__.py` (build()) first; there is no CPU fallback")
    L = ctypes.CDLL(p)
    vp, ci, cf, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t
    ip = ctypes.POINTER(ci)
    L.orbx_create.argtypes = [ctypes.POINTER(vp), ci, ci, cf, ci, ci, ci, ci, ci, ci]
    L.orbx_destroy.argtypes = [vp]
    L.orbx_destroy.restype = None
    L.orbx_strerror.argtypes = [ci]
    L.orbx_strerror.restype = ctypes.c_char_p
    L.orbx_last_cuda_error.argtypes = [vp]
    L.orbx_last_cuda_error.restype = ctypes.c_char_p
    L.orbx_get_levels.argtypes = [vp]
    L.orbx_get_scale_factor.argtypes = [vp]
    L.orbx_get_scale_factor.restype = cf
    L.orbx_get_tables.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orbx_max_keypoints.argtypes = [vp]
    L.orbx_extract.argtypes = [vp, vp, ci, ci, sz, ci, vp, vp, ci, ip]
    L.orbx_extract_batch.argtypes = [vp, vp, ci, ci, ci, sz, sz, ci, vp, vp, ci, vp]
    L.orbx_pyramid_level.argtypes = [vp, ci, ci, ci, ci, vp, sz, ip, ip]
    L.orbx_level_size.argtypes = [vp, ci, ci, ci, ip, ip]
    L.orbx_sync.argtypes = [vp]
    L.orbx_set_chunk.argtypes = [vp, ci]
    L.orbx_stream.argtypes = [vp]
    L.orbx_stream.restype = vp
    L.orbx_set_profiling.argtypes = [vp, ci]
    L.orbx_stage_name.argtypes = [ci]
    L.orbx_stage_name.restype = ctypes.c_char_p
    L.orbx_stage_stats.argtypes = [vp, ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong)]
    L.orbx_reset_stats.argtypes = [vp]
    L.orbx_kernel_launches.argtypes = [vp]
    L.orbx_kernel_launches.restype = ctypes.c_longlong
    L.orbx_descriptor_distance.argtypes = [vp, vp, vp, ci, ci, vp]
    L.orbx_cvt_gray.argtypes = [vp, vp, ci, ci, sz, ci, ci, vp, sz]
    L.orbx_undistort_keypoints.argtypes = [vp, vp, ci, cf, cf, cf, cf, vp, ci, ci, vp]
    L.orbx_hamming_best2.argtypes = [vp, vp, ci, vp, ci, vp, vp, vp, vp, ci, vp, ci, vp, vp, vp]
    L.orbx_pyramid_bordered.argtypes = [vp, ci, ci, ci, vp, sz, vp, vp]
    L.orbx_create_matcher.argtypes = [ctypes.POINTER(vp), ci]
    L.orbx_hamming_best2_static.argtypes = L.orbx_hamming_best2.argtypes      # the no-allocation variant (include/orbx.h)
    L.orbx_test_sincosf.argtypes = [ctypes.c_uint32, ctypes.c_uint32, vp, vp]
    L.orbx_test_fast_atan2.argtypes = [vp, ci, vp]
    L.orbx_test_popc_peak.argtypes = [ci, ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    _lib = L
    return L


def stage_names():
    L = lib()
    return [L.orbx_stage_name(i).decode() for i in range(L.orbx_stage_count())]


def _np(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _ptr(t):
    """Device/host pointer of a numpy array, a torch tensor, or None."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return t.ctypes.data_as(ctypes.c_void_p)
    return ctypes.c_void_p(t.data_ptr())


class ORBextractor:
    """Mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:75-216).

    ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST); calling the object with an
    8-bit single-channel image returns (keypoints, descriptors) exactly like operator()
    (src/ORBextractor.cc:1543-1658): keypoints as a KP_DTYPE array (cv::KeyPoint layout), descriptors
    N x 32 uint8.  `mask` is accepted and ignored, like the reference (:1543).
    """

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0, max_w=1920, max_h=1080,
                 max_batch=1):
        self._h = ctypes.c_void_p()
        self._lib = lib()
        rc = self._lib.orbx_create(ctypes.byref(self._h), device, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST,
                                   max_w, max_h, max_batch)
        if rc:
            self._h = ctypes.c_void_p()
            raise OrbxError(f"orbx_create: {self._lib.orbx_strerror(rc).decode()} ({rc})")
        self.nfeatures, self.nlevels, self.max_batch, self.device = nfeatures, nlevels, max_batch, device
        self.mvImagePyramid = _PyramidView(self)

    def _check(self, rc, what):
        if rc:
            raise OrbxError(f"{what}: {self._lib.orbx_strerror(rc).decode()} ({rc}) "
                            f"{self._lib.orbx_last_cuda_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.orbx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- getters (include/ORBextractor.h:115-149) ----
    def GetLevels(self):
        return self._lib.orbx_get_levels(self._h)

    def GetScaleFactor(self):
        return self._lib.orbx_get_scale_factor(self._h)

    def _tables(self):
        n = self.nlevels
        arrs = [np.zeros(n, np.float32) for _ in range(4)] + [np.zeros(n, np.int32)]
        self._check(self._lib.orbx_get_tables(self._h, *[_np(a) for a in arrs]), "orbx_get_tables")
        return arrs

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        return self._tables()[4]

    def max_keypoints(self):
        return self._lib.orbx_max_keypoints(self._h)

    # ---- operator() ----
    def __call__(self, image, mask=None):
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 expected (src/ORBextractor.cc:1553)"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        self._last_wh = (w, h)
        cap = self.max_keypoints()
        kps = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = ctypes.c_int(0)
        rc = self._lib.orbx_extract(self._h, _np(image), w, h, image.strides[0], 0, _np(kps), _np(desc), cap, ctypes.byref(n))
        self._check(rc, "orbx_extract")
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images):
        """images: (B, H, W) uint8 numpy array (host). Returns list of (kps, desc) per frame."""
        assert images.dtype == np.uint8 and images.ndim == 3
        images = np.ascontiguousarray(images)
        B, h, w = images.shape
        self._last_wh = (w, h)
        cap = self.max_keypoints()
        kps = np.empty((B, cap), KP_DTYPE)
        desc = np.empty((B, cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        rc = self._lib.orbx_extract_batch(self._h, _np(images), B, w, h, images.strides[1], images.strides[0], 0,
                                          _np(kps), _np(desc), cap, _np(n))
        self._check(rc, "orbx_extract_batch")
        return [(kps[b, :n[b]].copy(), desc[b, :n[b]].copy()) for b in range(B)]

    def extract_batch_device(self, images, kps, desc, n_out, in_device=True, dense=False):
        """Asynchronous device-resident path used by bench.py.  images: torch uint8 (B,H,W) on the ctx device
        (or pinned host tensor with in_device=False); kps: torch uint8 (B,cap,28); desc: torch uint8 (B,cap,32);
        n_out: torch int32 (B,), all on the device.  Returns immediately; call sync().  dense=True (ORBX_OUT_DENSE): frame b's
        records start at row sum(n_out[:b]) of the flattened (B*cap, ...) buffers instead of row b*cap."""
        B, h, w = images.shape
        self._last_wh = (w, h)
        cap = kps.shape[1]
        flags = OUT_DEVICE | (IN_DEVICE if in_device else 0) | (4 if dense else 0)
        rc = self._lib.orbx_extract_batch(self._h, _ptr(images), B, w, h, images.stride(1), images.stride(0), flags,
                                          _ptr(kps), _ptr(desc), cap, _ptr(n_out))
        self._check(rc, "orbx_extract_batch(device)")

    def set_chunk(self, frames):
        """Frames per pipeline chunk of extract_batch (0 = whole batch at once)."""
        self._check(self._lib.orbx_set_chunk(self._h, int(frames)), "orbx_set_chunk")

    def sync(self):
        self._check(self._lib.orbx_sync(self._h), "orbx_sync")

    def pyramid_level(self, level, frame=0, border=0):
        w, h = ctypes.c_int(), ctypes.c_int()
        self._check(self._lib.orbx_level_size(self._h, level, *self._last_wh, ctypes.byref(w), ctypes.byref(h)), "orbx_level_size")
        out = np.empty((h.value + 2 * border, w.value + 2 * border), np.uint8)
        self._check(self._lib.orbx_pyramid_level(self._h, frame, level, border, 0, _np(out), out.strides[0], ctypes.byref(w),
                                                 ctypes.byref(h)), "orbx_pyramid_level")
        return out

    def pyramid_bordered(self, frame=0, border=19):
        """All levels of one frame of the last call with their reflect-101 frame (mvImagePyramid's memory), one device pass:
        list of (h + 2 border, w + 2 border) uint8 views into one block."""
        n = self.nlevels
        off = (ctypes.c_size_t * n)()
        total = ctypes.c_size_t()
        self._check(self._lib.orbx_pyramid_bordered(self._h, frame, border, 0, None, 0, off, ctypes.byref(total)), "orbx_pyramid_bordered")
        block = np.empty(total.value, np.uint8)
        self._check(self._lib.orbx_pyramid_bordered(self._h, frame, border, 0, _np(block), block.size, off, ctypes.byref(total)),
                    "orbx_pyramid_bordered")
        out = []
        for l in range(n):
            w, h = ctypes.c_int(), ctypes.c_int()
            self._check(self._lib.orbx_level_size(self._h, l, *self._last_wh, ctypes.byref(w), ctypes.byref(h)), "orbx_level_size")
            ow, oh = w.value + 2 * border, h.value + 2 * border
            out.append(block[off[l]:off[l] + ow * oh].reshape(oh, ow))
        return out

    # ---- instrumentation ----
    def set_profiling(self, on):
        self._check(self._lib.orbx_set_profiling(self._h, int(on)), "orbx_set_profiling")

    def reset_stats(self):
        self._check(self._lib.orbx_reset_stats(self._h), "orbx_reset_stats")

    def stage_stats(self):
        out = {}
        for i, name in enumerate(stage_names()):
            ms, n = ctypes.c_double(), ctypes.c_longlong()
            self._check(self._lib.orbx_stage_stats(self._h, i, ctypes.byref(ms), ctypes.byref(n)), "orbx_stage_stats")
            out[name] = (ms.value, n.value)
        return out

    def kernel_launches(self):
        return self._lib.orbx_kernel_launches(self._h)

    def stream(self):
        return self._lib.orbx_stream(self._h)


class _PyramidView:
    """mvImagePyramid (include/ORBextractor.h:158): lazily downloaded level images of the last frame."""

    def __init__(self, ex):
        self._ex = ex

    def __len__(self):
        return self._ex.nlevels

    def __getitem__(self, level):
        return self._ex.pyramid_level(level)


FMT_RGB, FMT_BGR, FMT_RGBA, FMT_BGRA = 0, 1, 2, 3


def cvt_gray(ex, img, fmt=FMT_RGB):
    """cv::cvtColor(img, gray, {RGB,BGR,RGBA,BGRA}2GRAY) as called by Tracking::GrabImage* (src/Tracking.cc:252-278) on the GPU.
    img: (h, w, 3|4) uint8 (any row stride); returns (h, w) uint8."""
    img = np.asarray(img, np.uint8)
    h, w, ch = img.shape
    assert ch == (4 if fmt & 2 else 3) and img.strides[2] == 1 and img.strides[1] == ch
    out = np.empty((h, w), np.uint8)
    ex._check(ex._lib.orbx_cvt_gray(ex._h, img.ctypes.data, w, h, img.strides[0], fmt, 0, _np(out), w), "orbx_cvt_gray")
    return out


def undistort_keypoints(ex, kps, K, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:725-781) on the GPU; K = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3])."""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    d = np.ascontiguousarray(dist, np.float32)
    out = np.empty(max(len(kps), 1), KP_DTYPE)
    ex._check(ex._lib.orbx_undistort_keypoints(ex._h, _np(kps), len(kps), *[float(np.float32(x)) for x in K], _np(d), len(d), 0, _np(out)),
              "orbx_undistort_keypoints")
    return out[:len(kps)]


def descriptor_distance(ex, a, b):
    """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2123-2143) for n descriptor pairs on the GPU."""
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.empty(len(a), np.int32)
    ex._check(ex._lib.orbx_descriptor_distance(ex._h, _np(a), _np(b), len(a), 0, _np(out)), "orbx_descriptor_distance")
    return out


def hamming_best2(ex, q, t, q_seg=None, q_idx=None, t_seg=None, t_idx=None, t_valid=None):
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
    t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    if q_seg is None:
        q_seg = np.array([0, len(q)], np.int32)
        t_seg = np.array([0, len(t)], np.int32)
    q_seg = np.ascontiguousarray(q_seg, np.int32)
    t_seg = np.ascontiguousarray(t_seg, np.int32)
    q_idx = None if q_idx is None else np.ascontiguousarray(q_idx, np.int32)
    t_idx = None if t_idx is None else np.ascontiguousarray(t_idx, np.int32)
    t_valid = None if t_valid is None else np.ascontiguousarray(t_valid, np.uint8)
    nrow = len(q) if q_idx is None else len(q_idx)
    bi = np.empty(nrow, np.int32)
    bd = np.empty(nrow, np.int32)
    sd = np.empty(nrow, np.int32)
    rc = ex._lib.orbx_hamming_best2(ex._h, _np(q), len(q), _np(t), len(t), _np(q_seg), _np(q_idx), _np(t_seg), _np(t_idx),
                                    len(q_seg) - 1, _np(t_valid), 0, _np(bi), _np(bd), _np(sd))
    ex._check(rc, "orbx_hamming_best2")
    return bi, bd, sd
