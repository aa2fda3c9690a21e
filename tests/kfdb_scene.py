"""Scripts of KeyFrameDatabase operations (add / erase / set covisible / loop query / relocalisation query) and the adapters
that replay them on the oracle (oracle/orb_kfdb.c), on the unmodified reference (oracle/_ref) and on the device
(orbx_kfdb_*).  Every backend returns, per query, the ordered list of candidate keyframe ids."""
import ctypes
import numpy as np

I32P = ctypes.POINTER(ctypes.c_int32)
F64P = ctypes.POINTER(ctypes.c_double)


def _i32(a):
    a = np.ascontiguousarray(a, np.int32)
    return a, a.ctypes.data_as(I32P)


def _f64(a):
    a = np.ascontiguousarray(a, np.float64)
    return a, a.ctypes.data_as(F64P)


def bow_vector(rng, nwords, centre, n, spread):
    """A sorted BowVector of about n distinct words around `centre` (places overlap when their centres are close), L1-normalised
    tf-idf-like weights."""
    w = np.unique(np.clip((centre + rng.normal(0, spread, n)).astype(np.int64), 0, nwords - 1)).astype(np.int32)
    v = rng.random(len(w)) * rng.integers(1, 4, len(w)) + 0.05
    return w, (v / v.sum()).astype(np.float64)


def make_script(seed, nkf=300, nwords=20000, nqueries=40, words_per_kf=220, places=12, erase_frac=0.05, spread=260.0):
    """Keyframes visit `places` (word clusters) in a loop, so that later keyframes share many words with earlier ones.
    Returns a list of operations; queries are interleaved with the insertions the way LoopClosing / Tracking issue them."""
    rng = np.random.default_rng(seed)
    centres = rng.integers(1000, nwords - 1000, places)
    ops, bows, alive = [], {}, []
    q_every = max(1, nkf // max(1, nqueries))
    frame_id = 5
    for k in range(nkf):
        c = centres[(k // 6) % places] + rng.normal(0, 60)
        w, v = bow_vector(rng, nwords, c, words_per_kf, spread)
        bows[k] = (w, v)
        ops.append(("new", k, w, v))
        if k >= 3 and k % q_every == 0:                   # loop query BEFORE the keyframe enters the database (LoopClosing.cc order)
            conn = [int(x) for x in rng.choice(alive, min(len(alive), int(rng.integers(0, 12))), replace=False)] if alive else []
            conn += [k - 1, k - 2]
            min_score = float(rng.choice([0.0, 0.01, 0.03, 0.08]))
            ops.append(("loop", k, np.array(conn, np.int32), min_score))
        ops.append(("add", k))
        alive.append(k)
        # covisibility: mostly recent keyframes and a few old ones, some not (or no longer) in the database
        pool = [x for x in range(max(0, k - 25), k)] + [int(x) for x in rng.integers(0, k + 1, 4)]
        nb = [int(x) for x in rng.choice(pool, min(len(pool), int(rng.integers(0, 14))), replace=False)] if pool else []
        ops.append(("covis", k, np.array(nb, np.int32)))
        if k % 9 == 4 and k > 10:                          # refresh an older keyframe's list, like UpdateConnections does
            j = int(rng.integers(0, k))
            nb = [int(x) for x in rng.integers(0, k + 1, int(rng.integers(0, 12)))]
            ops.append(("covis", j, np.array(nb, np.int32)))
        if k > 20 and rng.random() < erase_frac:
            j = int(rng.choice(alive))
            alive.remove(j)
            ops.append(("erase", j))
        if k >= 3 and k % q_every == q_every // 2:         # relocalisation with a frame near a random place
            c = centres[int(rng.integers(0, places))] + rng.normal(0, 80)
            w, v = bow_vector(rng, nwords, c, words_per_kf, spread)
            frame_id += int(rng.integers(1, 4))
            ops.append(("reloc", frame_id, w, v))
            if rng.random() < 0.3:                         # the same place again: scores of the previous query are still in the keyframes
                w2, v2 = bow_vector(rng, nwords, c + 400, words_per_kf // 3, spread)
                frame_id += 1
                ops.append(("reloc", frame_id, w2, v2))
    return ops, nwords


class OracleBackend:
    def __init__(self, nwords):
        import oracle_lib as O
        self.L = O.lib()
        L = self.L
        L.orc_kfdb_create.restype = ctypes.c_void_p
        L.orc_kfdb_create.argtypes = [ctypes.c_int]
        L.orc_kfdb_destroy.argtypes = [ctypes.c_void_p]
        L.orc_kfdb_clear.argtypes = [ctypes.c_void_p]
        L.orc_kfdb_new_keyframe.argtypes = [ctypes.c_void_p, ctypes.c_int, I32P, F64P, ctypes.c_int]
        L.orc_kfdb_add.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_kfdb_erase.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_kfdb_set_covisible.argtypes = [ctypes.c_void_p, ctypes.c_int, I32P, ctypes.c_int]
        L.orc_kfdb_detect_loop.argtypes = [ctypes.c_void_p, ctypes.c_int, I32P, ctypes.c_int, ctypes.c_float, I32P, ctypes.c_int]
        L.orc_kfdb_detect_reloc.argtypes = [ctypes.c_void_p, ctypes.c_longlong, I32P, F64P, ctypes.c_int, I32P, ctypes.c_int]
        self.h = L.orc_kfdb_create(nwords)
        self.p = "orc_kfdb_"

    def _f(self, name):
        return getattr(self.L, self.p + name)

    def new(self, k, w, v):
        w, wp = _i32(w); v, vp = _f64(v)
        assert self._f("new_keyframe")(self.h, k, wp, vp, len(w)) == 0

    def add(self, k):
        assert self._f("add")(self.h, k) == 0

    def erase(self, k):
        assert self._f("erase")(self.h, k) == 0

    def covis(self, k, ids):
        a, ap = _i32(ids)
        assert self._f("set_covisible")(self.h, k, ap, len(a)) == 0

    def loop(self, k, conn, min_score, cap=4096):
        a, ap = _i32(conn)
        out = np.zeros(cap, np.int32)
        n = self._f("detect_loop")(self.h, k, ap, len(a), min_score, out.ctypes.data_as(I32P), cap)
        assert 0 <= n <= cap, n
        return out[:n].tolist()

    def reloc(self, fid, w, v, cap=4096):
        w, wp = _i32(w); v, vp = _f64(v)
        out = np.zeros(cap, np.int32)
        n = self._f("detect_reloc")(self.h, fid, wp, vp, len(w), out.ctypes.data_as(I32P), cap)
        assert 0 <= n <= cap, n
        return out[:n].tolist()

    def clear(self):
        self._f("clear")(self.h)

    def close(self):
        self._f("destroy")(self.h)


class RefBackend(OracleBackend):
    """The reference's own KeyFrameDatabase (oracle/_ref/liborb_ref.so: ref_kfdb_*)."""

    def __init__(self, nwords):
        import ref_lib as R
        self.L = R.lib()
        L = self.L
        L.ref_kfdb_create.restype = ctypes.c_void_p
        L.ref_kfdb_create.argtypes = [ctypes.c_int]
        L.ref_kfdb_destroy.argtypes = [ctypes.c_void_p]
        L.ref_kfdb_clear.argtypes = [ctypes.c_void_p]
        L.ref_kfdb_new_keyframe.argtypes = [ctypes.c_void_p, ctypes.c_int, I32P, F64P, ctypes.c_int]
        L.ref_kfdb_add.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_kfdb_erase.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_kfdb_set_covisible.argtypes = [ctypes.c_void_p, ctypes.c_int, I32P, ctypes.c_int]
        L.ref_kfdb_detect_loop.argtypes = [ctypes.c_void_p, ctypes.c_int, I32P, ctypes.c_int, ctypes.c_float, I32P, ctypes.c_int]
        L.ref_kfdb_detect_reloc.argtypes = [ctypes.c_void_p, ctypes.c_longlong, I32P, F64P, ctypes.c_int, I32P, ctypes.c_int]
        self.h = L.ref_kfdb_create(nwords)
        self.p = "ref_kfdb_"


class GpuBackend:
    """orbx_kfdb_* through the package's KeyFrameDatabase; keyframe objects (id -> BowVector) are kept here, the database
    only sees them at add()."""

    def __init__(self, nwords, ctx):
        import importlib
        P = importlib.import_module("orb-slam2-noted_b200")
        self.db = P.KeyFrameDatabase(ctx, nwords)
        self.bow, self.covis_of, self.in_db = {}, {}, set()

    def new(self, k, w, v):
        self.bow[k] = (w, v)

    def add(self, k):
        self.db.add(k, *self.bow[k])
        self.in_db.add(k)
        if k in self.covis_of:                           # the list was set before the keyframe entered the database
            self.db.set_covisible(k, self.covis_of[k])

    def erase(self, k):
        self.db.erase(k)
        self.in_db.discard(k)

    def covis(self, k, ids):                             # a property of the keyframe; the database needs it for the ones it holds
        self.covis_of[k] = ids
        if k in self.in_db:
            self.db.set_covisible(k, ids)

    def loop(self, k, conn, min_score, cap=4096):
        return self.db.DetectLoopCandidates(k, *self.bow[k], conn, min_score, cap)

    def reloc(self, fid, w, v, cap=4096):
        return self.db.DetectRelocalizationCandidates(fid, w, v, cap)

    def clear(self):
        self.db.clear()
        self.in_db.clear()

    def close(self):
        self.db.close()


def replay(ops, backend):
    """Returns [(kind, id, [candidate ids])] for the queries of the script."""
    res = []
    for op in ops:
        kind = op[0]
        if kind == "new":
            backend.new(op[1], op[2], op[3])
        elif kind == "add":
            backend.add(op[1])
        elif kind == "erase":
            backend.erase(op[1])
        elif kind == "covis":
            backend.covis(op[1], op[2])
        elif kind == "loop":
            res.append(("loop", op[1], backend.loop(op[1], op[2], op[3])))
        elif kind == "reloc":
            res.append(("reloc", op[1], backend.reloc(op[1], op[2], op[3])))
        elif kind == "clear":
            backend.clear()
    return res
