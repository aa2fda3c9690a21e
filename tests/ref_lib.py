"""ctypes binding of oracle/_ref/liborb_ref.so: the UNMODIFIED reference sources compiled against oracle/cvshim
(TEST INFRASTRUCTURE: the pin the oracle and the GPU path are checked against, never the product).
The library is built by `make -C oracle ref` where /root/reference exists (this container); on the GPU box the prebuilt
file travels with the snapshot.  available() is False when neither holds, and the tests that need it skip."""
import ctypes
import os
import subprocess
import numpy as np
import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
SO = os.path.join(ODIR, "_ref", "liborb_ref.so")
REF = "/root/reference"
_lib = None


def available():
    return os.path.exists(SO) or os.path.exists(os.path.join(REF, "src", "ORBextractor.cc"))


def lib():
    global _lib
    if _lib is None:
        if os.path.exists(os.path.join(REF, "src", "ORBextractor.cc")):
            subprocess.run(["make", "-C", ODIR, "-s", "ref"], check=True)
        L = ctypes.CDLL(SO)
        vp, ci, cf, cd = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_double
        L.ref_extract.argtypes = [ci, cf, ci, ci, ci, vp, ci, ci, ctypes.c_size_t, vp, vp, ci, vp, vp]
        L.ref_extractor_tables.argtypes = [ci, cf, ci, ci, ci, vp, vp, vp, vp, vp, vp]
        L.ref_extractor_tables.restype = None
        L.ref_distribute_octree.argtypes = [vp, ci, ci, ci, ci, ci, ci, ci, vp, ci]
        L.ref_extract_throughput.argtypes = [ci, cf, ci, ci, ci, vp, ci, ci, ci, ci, ci, vp]
        L.ref_extract_throughput.restype = cd
        L.ref_stereo_frame.argtypes = [ci, cf, ci, ci, ci, vp, vp, ci, ci, cf, cf, cf, cf, vp, ci, cf, cf, vp, vp, vp, vp, vp, vp, ci, vp, vp, vp]
        L.ref_vocab_load_text.argtypes = [ctypes.c_char_p]
        L.ref_vocab_load_text.restype = vp
        L.ref_vocab_save_text.argtypes = [vp, ctypes.c_char_p]
        L.ref_vocab_save_text.restype = None
        L.ref_vocab_free.argtypes = [vp]
        L.ref_vocab_free.restype = None
        L.ref_vocab_info.argtypes = [vp] * 7
        L.ref_vocab_info.restype = None
        L.ref_vocab_transform_each.argtypes = [vp, vp, ci, ci, vp, vp, vp]
        L.ref_vocab_transform_each.restype = None
        L.ref_vocab_transform.argtypes = [vp, vp, ci, ci, vp, vp, vp, vp, vp, vp]
        L.ref_vocab_score.argtypes = [vp, vp, vp, ci, vp, vp, ci]
        L.ref_vocab_score.restype = cd
        L.ref_descriptor_distance.argtypes = [vp, vp]
        L.ref_hamming_throughput.argtypes = [vp, ci, vp, ci, ci, vp, vp, vp]
        L.ref_hamming_throughput.restype = cd
        L.ref_shim_algebra.argtypes = [ci, vp, ci, ci, vp, ci, ci, vp, vp]
        L.ref_shim_norm.argtypes = [vp, ci]
        L.ref_shim_norm.restype = cd
        L.ref_shim_dot.argtypes = [vp, vp, ci]
        L.ref_shim_dot.restype = cd
        L.ref_run_matcher_scene.argtypes = [ctypes.c_char_p, ctypes.c_char_p]
        L.ref_set_allocator.argtypes = [ci]
        assert L.ref_set_allocator(1) == 0      # canonical (monotonic) model by default; set_allocator(0) = the real heap
        L.ref_set_allocator.argtypes = [ci]
        assert L.ref_set_allocator(1) == 0      # canonical (monotonic) model by default; set_allocator(0) = the real heap
        _lib = L
    return _lib


def set_allocator(mode):
    """0: the process's real heap (glibc); 1: monotonic small-object addresses (address order == creation order)."""
    assert lib().ref_set_allocator(int(mode)) == 0


def set_allocator(mode):
    """0: the process's real heap (glibc); 1: monotonic small-object addresses (address order == creation order)."""
    assert lib().ref_set_allocator(int(mode)) == 0


_p = lambda a: None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def extract(img, nfeatures=1000, scale=1.2, nlevels=8, ini=20, mn=7, want_pyr=False, want_bordered=False):
    """ORBextractor::operator() of the reference.  Returns (kps, desc[, pyr][, bordered pyr])."""
    img = np.ascontiguousarray(img)
    h, w = img.shape
    cap = nfeatures + 64 * nlevels + 1024
    kps = np.empty(cap, O.KP_DTYPE)
    desc = np.empty((cap, 32), np.uint8)
    sizes = O.level_sizes(O.params(nfeatures, scale, nlevels, ini, mn), w, h)
    pyr = np.empty(sum(a * b for a, b in sizes), np.uint8) if want_pyr else None
    pb = np.empty(sum((a + 38) * (b + 38) for a, b in sizes), np.uint8) if want_bordered else None
    n = lib().ref_extract(nfeatures, scale, nlevels, ini, mn, _p(img), w, h, img.strides[0], _p(kps), _p(desc), cap, _p(pyr), _p(pb))
    assert 0 <= n <= cap, n
    out = [kps[:n].copy(), desc[:n].copy()]
    if want_pyr:
        out.append(pyr)
    if want_bordered:
        out.append(pb)
    return tuple(out)


def distribute(kps, minx, maxx, miny, maxy, N, level=0):
    kps = np.ascontiguousarray(kps)
    out = np.empty(max(len(kps), 1), O.KP_DTYPE)
    n = lib().ref_distribute_octree(_p(kps), len(kps), minx, maxx, miny, maxy, N, level, _p(out), len(out))
    assert n >= 0
    return out[:n].copy()
