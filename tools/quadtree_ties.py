#!/usr/bin/env python3
"""How often does the reference's heap-address tie-break (sort(pair<int, ExtractorNode*>), src/ORBextractor.cc:935) order
equal-size quadtree nodes differently from the canonical rule (address order == creation order) that the oracle and the
GPU kernel implement?  Runs the UNMODIFIED reference (oracle/_ref) twice per synthetic frame -- with the process's real
glibc heap and with monotonic node addresses -- and compares both with the oracle.  Build container only.
  python tools/quadtree_ties.py [--frames 1000] [--out profiles/r2_quadtree_ties.json]"""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth, oracle_lib as O, ref_lib as R

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=1000)
ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_quadtree_ties.json"))
a = ap.parse_args()
shapes = [("640x480/1000", 640, 480, 1000), ("752x480/1200", 752, 480, 1200), ("1241x376/2000", 1241, 376, 2000)]
res = {}
for name, w, h, nf in shapes:
    p = O.params(nf)
    st = dict(frames=0, levels=0, mono_frames_equal=0, heap_frames_equal=0, heap_levels_differ=0, heap_kps_total=0,
              heap_kps_not_in_canonical=0, canonical_kps_not_in_heap=0, heap_count_delta_max=0)
    n = a.frames if w == 640 else max(50, a.frames // 5)
    for seed in range(n):
        img = synth.frame(1000 + seed, w, h)
        ok, od = O.extract(p, img)
        R.set_allocator(1); mk, md = R.extract(img, nf)
        R.set_allocator(0); hk, hd = R.extract(img, nf)
        st["frames"] += 1; st["levels"] += 8
        st["mono_frames_equal"] += int(mk.tobytes() == ok.tobytes() and np.array_equal(md, od))
        st["heap_frames_equal"] += int(hk.tobytes() == ok.tobytes() and np.array_equal(hd, od))
        for l in range(8):
            x, y = ok[ok["octave"] == l], hk[hk["octave"] == l]
            if x.tobytes() != y.tobytes():
                st["heap_levels_differ"] += 1
        so, sh = set(map(bytes, ok)), set(map(bytes, hk))
        st["heap_kps_total"] += len(hk)
        st["heap_kps_not_in_canonical"] += len(sh - so)
        st["canonical_kps_not_in_heap"] += len(so - sh)
        st["heap_count_delta_max"] = max(st["heap_count_delta_max"], abs(len(hk) - len(ok)))
    R.set_allocator(1)
    res[name] = st
    print(name, st, flush=True)
json.dump(dict(what="reference (oracle/_ref, unmodified) vs canonical tie-break; synth.frame seeds 1000..", results=res), open(a.out, "w"), indent=1)
