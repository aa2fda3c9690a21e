"""KeyFrameDatabase queries: device (orbx_kfdb_*) against the unmodified reference class on the host (oracle/_ref), same
database, same queries, results compared.  Sized like a long ORB-SLAM2 session: k=10 L=6 vocabulary (10^6 words), a few
thousand keyframes of ~900 words each.   python tools/bench_kfdb.py [keyframes] [queries]  -> one JSON line"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import kfdb_scene as K
import ref_lib as R


def measure(nkf=4000, nq=200, ex=None):
    P = importlib.import_module("orb-slam2-noted_b200")
    nwords, per_kf, places = 1000000, 900, 60
    rng = np.random.default_rng(1)
    centres = rng.integers(20000, nwords - 20000, places)
    bows = [K.bow_vector(rng, nwords, centres[(k // 8) % places] + rng.normal(0, 300), per_kf, 2500.0) for k in range(nkf)]
    covis = [np.array([x for x in range(max(0, k - 10), k)][::-1], np.int32) for k in range(nkf)]
    queries = [K.bow_vector(rng, nwords, centres[int(rng.integers(0, places))] + rng.normal(0, 300), per_kf, 2500.0) for _ in range(nq)]
    if ex is None:
        ex = P.ORBextractor(500, 1.2, 8, 20, 7, max_w=64, max_h=64, max_batch=1)
    g = K.GpuBackend(nwords, ex)
    t0 = time.perf_counter()
    for k in range(nkf):
        g.new(k, *bows[k]); g.add(k); g.covis(k, covis[k])
    ex.sync()
    t_add = (time.perf_counter() - t0) / nkf
    ids = list(range(10, 10 + nq))
    g.db.detect_reloc_batch(ids[:8], queries[:8], cap=256)                   # warm-up (allocations)
    t0 = time.perf_counter()
    one = [g.db.DetectRelocalizationCandidates(1000 + i, *queries[i], cap=256) for i in range(nq)]
    t_one = (time.perf_counter() - t0) / nq
    t0 = time.perf_counter()
    batch = g.db.detect_reloc_batch([5000 + i for i in range(nq)], queries, cap=256)
    t_batch = (time.perf_counter() - t0) / nq
    t0 = time.perf_counter()
    loops = [g.db.DetectLoopCandidates(20000 + i, *queries[i], np.arange(i, i + 15, dtype=np.int32), 0.01, cap=256) for i in range(nq)]
    t_loop = (time.perf_counter() - t0) / nq
    out = {"keyframes": nkf, "words_per_keyframe": float(np.mean([len(b[0]) for b in bows])), "vocabulary_words": nwords, "queries": nq,
           "gpu_add_ms": t_add * 1e3, "gpu_reloc_ms_per_query_single_calls": t_one * 1e3, "gpu_reloc_ms_per_query_batched": t_batch * 1e3,
           "gpu_loop_ms_per_query": t_loop * 1e3, "mean_candidates": float(np.mean([len(x) for x in one]))}
    if os.path.exists(R.SO):
        r = K.RefBackend(nwords)
        for k in range(nkf):
            r.new(k, *bows[k]); r.add(k); r.covis(k, covis[k])
        t0 = time.perf_counter()
        ref_one = [r.reloc(1000 + i, *queries[i], cap=256) for i in range(nq)]
        t_ref = (time.perf_counter() - t0) / nq
        ref_batch = [r.reloc(5000 + i, *queries[i], cap=256) for i in range(nq)]
        for i in range(nq):
            r.new(20000 + i, *queries[i])
        t0 = time.perf_counter()
        ref_loops = [r.loop(20000 + i, np.arange(i, i + 15, dtype=np.int32), 0.01, cap=256) for i in range(nq)]
        t_refl = (time.perf_counter() - t0) / nq
        out.update({"cpu_reference_reloc_ms_per_query": t_ref * 1e3, "cpu_reference_loop_ms_per_query": t_refl * 1e3,
                    "identical_to_reference": one == ref_one and batch == ref_batch and loops == ref_loops,
                    "cpu": "oracle/_ref: unmodified src/KeyFrameDatabase.cc, 1 thread (the reference holds a mutex for the whole walk)"})
        r.close()
    g.close()
    return out


if __name__ == "__main__":
    print(json.dumps(measure(int(sys.argv[1]) if len(sys.argv) > 1 else 4000, int(sys.argv[2]) if len(sys.argv) > 2 else 200)))
