#!/usr/bin/env python3
"""SearchByBoW(KeyFrame, Frame) wall time on the config-4 shape (10 k x 10 k descriptors per keyframe pair), brute-force
mode (levelsup >= L: one node) and BoW mode, through the host-buffer C ABI; also prints the per-stage device time."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth


def run(n=10000, reps=5):
    P = importlib.import_module("orb-slam2-noted_b200")
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)
    v = synth.vocabulary(0, 10, 3)
    voc = P.Vocabulary(ex, v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    rng = np.random.default_rng(0)
    d1 = synth.descriptors(1, n)
    d2 = np.concatenate([synth.descriptors(2, n * 3 // 4, 0.1, d1[: n * 3 // 4]), synth.descriptors(3, n - n * 3 // 4)])[rng.permutation(n)]
    k1, k2 = synth.keypoints(4, n, dtype=P.KP_DTYPE), synth.keypoints(5, n, dtype=P.KP_DTYPE)
    out = {}
    for name, levelsup in (("brute_force", 4), ("bow", 1)):
        sides = []
        for d, k in ((d1, k1), (d2, k2)):
            w, wt, nd = voc.transform(d, levelsup)
            nid, noff, feat = P.feature_vector(nd, wt)
            sides.append(P.FrameView(d, k, np.ones(n, np.uint8), None, nid, noff, feat))
        m = P.ORBmatcher(ex, 0.7, True)
        m.SearchByBoW(sides[0], sides[1])
        ex.set_profiling(True); ex.reset_stats()
        t0 = time.perf_counter()
        for _ in range(reps):
            match, nm = m.SearchByBoW(sides[0], sides[1])
        dt = (time.perf_counter() - t0) / reps
        st = {k: round(val[0] / reps, 3) for k, val in ex.stage_stats().items() if val[0] > 0}
        ex.set_profiling(False)
        pairs = sum(int(noff1[i + 1] - noff1[i]) for noff1 in [sides[0].node_off] for i in range(len(noff1) - 1)) if False else None
        # throughput with many pairs per call (orbx_match_bow_batch)
        npairs = 64 if name == "brute_force" else 256
        batch = [(sides[0], sides[1])] * npairs
        m.SearchByBoWBatch(batch[:2])
        t0 = time.perf_counter()
        res = m.SearchByBoWBatch(batch)
        dtb = time.perf_counter() - t0
        assert all(np.array_equal(r[0], match) for r in res[:3])
        out[name + "_batch"] = {"pairs": npairs, "ms_per_pair_wall": dtb * 1e3 / npairs,
                                "gmatch_per_s_effective": npairs * sum(int(sides[0].node_off[i + 1] - sides[0].node_off[i]) *
                                    int(sides[1].node_off[list(sides[1].node_id).index(nid) + 1] - sides[1].node_off[list(sides[1].node_id).index(nid)])
                                    for i, nid in enumerate(sides[0].node_id) if nid in set(sides[1].node_id)) / dtb / 1e9}
        out[name] = {"ms_per_pair_wall": dt * 1e3, "matches": int(nm), "nodes": int(len(sides[0].node_id)), "device_ms": st}
    return out


if __name__ == "__main__":
    print(json.dumps(run(int(sys.argv[1]) if len(sys.argv) > 1 else 10000), indent=1))
