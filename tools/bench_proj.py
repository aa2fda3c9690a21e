#!/usr/bin/env python3
"""Wall time of the tracking-thread matchers on a typical frame: SearchByProjection(Frame, MapPoints) with 2000 keypoints /
2000 projected points, and ComputeStereoMatches-free grid assignment; prints per-stage device time."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth, oracle_lib as O
import test_gpu_match as T


def run(n=2000, nq=2000, reps=20):
    P = importlib.import_module("orb-slam2-noted_b200")
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)
    keys, desc, uright, occupied, pts = T.proj_scene(P, 1, n, nq)
    scale = np.array(O.params(1000).scale[:8], np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    fv = P.FrameView(desc, keys, None, uright)
    out = {}
    for name, th in (("th1", 1.0), ("th15", 15.0)):
        q = np.zeros(nq, P.PROJ_QUERY_DTYPE)
        q["u"], q["v"], q["ur"], q["angle"], q["desc"] = pts["u"], pts["v"], pts["ur"], pts["angle"], pts["desc"]
        q["radius"] = (np.float32(4.0 * th) * scale[pts["level"]]).astype(np.float32)
        q["min_level"], q["max_level"] = pts["level"] - 1, pts["level"]
        m = P.ORBmatcher(ex, 0.8, True)
        m.SearchByProjection(fv, q, grid, 100, occupied, ratio_same_level=True)
        ex.set_profiling(True); ex.reset_stats()
        t0 = time.perf_counter()
        for _ in range(reps):
            got, ng = m.SearchByProjection(fv, q, grid, 100, occupied, ratio_same_level=True)
        dt = (time.perf_counter() - t0) / reps
        st = {k: round(v[0] / reps, 3) for k, v in ex.stage_stats().items() if v[0] > 0}
        ex.set_profiling(False)
        t0 = time.perf_counter()
        want, nw = O.search_by_projection(pts, keys, desc, uright, scale, grid, occupied, 4.0 * th / 4.0 if False else th, nnratio=0.8)
        cpu = time.perf_counter() - t0
        out[name] = {"ms_wall": dt * 1e3, "matches": int(ng), "device_ms": st, "oracle_cpu_ms_1thread": cpu * 1e3}
    return out


if __name__ == "__main__":
    print(json.dumps(run(), indent=1))
