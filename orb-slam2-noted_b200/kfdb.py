"""KeyFrameDatabase on the device: the host-side mirror of the reference's class (include/KeyFrameDatabase.h,
src/KeyFrameDatabase.cc) over orbx_kfdb_* (csrc/k_kfdb.cu).  Keyframes are named by their ids (KeyFrame::mnId)."""
import ctypes
import numpy as np
from .binding import lib, OrbxError

_I32P = ctypes.POINTER(ctypes.c_int32)
_F64P = ctypes.POINTER(ctypes.c_double)
_I64P = ctypes.POINTER(ctypes.c_longlong)


def _bind(L):
    if getattr(L, "_kfdb_bound", False):
        return
    vp, ci = ctypes.c_void_p, ctypes.c_int
    L.orbx_kfdb_create.argtypes = [vp, ci, ctypes.POINTER(vp)]
    L.orbx_kfdb_destroy.argtypes = [vp]
    L.orbx_kfdb_destroy.restype = None
    L.orbx_kfdb_size.argtypes = [vp]
    L.orbx_kfdb_add.argtypes = [vp, ci, _I32P, _F64P, ci]
    L.orbx_kfdb_erase.argtypes = [vp, ci]
    L.orbx_kfdb_clear.argtypes = [vp]
    L.orbx_kfdb_set_covisible.argtypes = [vp, ci, _I32P, ci]
    L.orbx_kfdb_detect_loop.argtypes = [vp, ctypes.c_longlong, _I32P, _F64P, ci, _I32P, ci, ctypes.c_float, _I32P, ci, ctypes.POINTER(ci)]
    L.orbx_kfdb_detect_reloc.argtypes = [vp, ci, _I64P, _I32P, _F64P, _I32P, _I32P, ci, _I32P]
    L._kfdb_bound = True


def _i32(a):
    a = np.ascontiguousarray(a, np.int32)
    return a, a.ctypes.data_as(_I32P)


def _f64(a):
    a = np.ascontiguousarray(a, np.float64)
    return a, a.ctypes.data_as(_F64P)


class KeyFrameDatabase:
    """`ctx` is anything that owns an orbx context (`._h`): an ORBextractor or a matcher context.  `nwords` = voc.size()."""

    def __init__(self, ctx, nwords):
        self._L = lib()
        _bind(self._L)
        self._ctx = ctx
        self._h = ctypes.c_void_p()
        self._chk(self._L.orbx_kfdb_create(ctx._h, int(nwords), ctypes.byref(self._h)), "orbx_kfdb_create")

    def _chk(self, rc, what):
        if rc < 0:
            raise OrbxError(f"{what}: {self._L.orbx_strerror(rc).decode()} ({rc}) {self._L.orbx_last_cuda_error(self._ctx._h).decode()}")
        return rc

    def __len__(self):
        return self._chk(self._L.orbx_kfdb_size(self._h), "orbx_kfdb_size")

    def add(self, kf_id, words, vals):
        """KeyFrameDatabase::add (src/KeyFrameDatabase.cc:50-56): the keyframe's BowVector as ascending word ids + weights."""
        w, wp = _i32(words); v, vp = _f64(vals)
        self._chk(self._L.orbx_kfdb_add(self._h, int(kf_id), wp, vp, len(w)), "orbx_kfdb_add")

    def erase(self, kf_id):
        self._chk(self._L.orbx_kfdb_erase(self._h, int(kf_id)), "orbx_kfdb_erase")

    def clear(self):
        self._chk(self._L.orbx_kfdb_clear(self._h), "orbx_kfdb_clear")

    def set_covisible(self, kf_id, ids):
        a, ap = _i32(ids)
        self._chk(self._L.orbx_kfdb_set_covisible(self._h, int(kf_id), ap, len(a)), "orbx_kfdb_set_covisible")

    def DetectLoopCandidates(self, query_id, words, vals, connected, min_score, cap=1024):
        """:113-270 -> candidate keyframe ids in the reference's order."""
        w, wp = _i32(words); v, vp = _f64(vals); cn, cp = _i32(connected)
        while True:
            out = np.zeros(cap, np.int32)
            n = ctypes.c_int()
            self._chk(self._L.orbx_kfdb_detect_loop(self._h, int(query_id), wp, vp, len(w), cp, len(cn), float(min_score),
                                                    out.ctypes.data_as(_I32P), cap, ctypes.byref(n)), "orbx_kfdb_detect_loop")
            if n.value <= cap:
                return out[:n.value].tolist()
            raise OrbxError(f"DetectLoopCandidates: {n.value} candidates, cap {cap} (the query state has advanced; pass a larger cap)")

    def DetectRelocalizationCandidates(self, frame_id, words, vals, cap=1024):
        return self.detect_reloc_batch([frame_id], [(words, vals)], cap)[0]

    def detect_reloc_batch(self, frame_ids, bows, cap=256):
        """:275-399 for several frames in one call (in order, one synchronisation) -> a list of candidate-id lists."""
        nq = len(frame_ids)
        seg = np.zeros(nq + 1, np.int32)
        for i, (w, _) in enumerate(bows):
            seg[i + 1] = seg[i] + len(w)
        words = np.concatenate([np.asarray(w, np.int32) for w, _ in bows]) if nq else np.zeros(0, np.int32)
        vals = np.concatenate([np.asarray(v, np.float64) for _, v in bows]) if nq else np.zeros(0, np.float64)
        w, wp = _i32(words); v, vp = _f64(vals); s, sp = _i32(seg)
        q = np.ascontiguousarray(frame_ids, np.int64)
        out = np.zeros((nq, cap), np.int32)
        n = np.zeros(nq, np.int32)
        self._chk(self._L.orbx_kfdb_detect_reloc(self._h, nq, q.ctypes.data_as(_I64P), wp, vp, sp, out.ctypes.data_as(_I32P), cap,
                                                 n.ctypes.data_as(_I32P)), "orbx_kfdb_detect_reloc")
        if nq and int(n.max()) > cap:
            raise OrbxError(f"DetectRelocalizationCandidates: {int(n.max())} candidates, cap {cap}")
        return [out[i, :n[i]].tolist() for i in range(nq)]

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbx_kfdb_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
