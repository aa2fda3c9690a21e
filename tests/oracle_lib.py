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
