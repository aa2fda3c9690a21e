#!/usr/bin/env python3
"""Hamming GMatch/s of the many-vs-many kernel (K6): `pairs` keyframe pairs of n x n descriptors, device-resident,
one orbx_hamming_best2 call (segments = pairs).  Prints GMatch/s and the fraction of the measured POPC peak."""
import ctypes, importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def run(n=10000, pairs=32, reps=5):
    import torch
    P = importlib.import_module("orb-slam2-noted_b200")
    L = P.lib()
    ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(0)
    q = torch.randint(0, 256, (pairs * n, 32), dtype=torch.uint8, generator=g).to(dev)
    t = torch.randint(0, 256, (pairs * n, 32), dtype=torch.uint8, generator=g).to(dev)
    seg = np.arange(pairs + 1, dtype=np.int32) * n
    bi = torch.empty(pairs * n, dtype=torch.int32, device=dev)
    bd = torch.empty_like(bi); sd = torch.empty_like(bi)
    vp = ctypes.c_void_p

    def call():
        rc = L.orbx_hamming_best2(ex._h, vp(q.data_ptr()), pairs * n, vp(t.data_ptr()), pairs * n, seg.ctypes.data_as(vp), None,
                                  seg.ctypes.data_as(vp), None, pairs, None, 3, vp(bi.data_ptr()), vp(bd.data_ptr()), vp(sd.data_ptr()))
        assert rc == 0, rc
    call(); ex.sync()
    ex.set_profiling(True); ex.reset_stats()
    for _ in range(reps):
        call()
    ex.sync()
    ms = ex.stage_stats()["hamming"][0] / reps
    peak = ctypes.c_double(); pms = ctypes.c_double()
    L.orbx_test_popc_peak(0, 1 << 16, ctypes.byref(peak), ctypes.byref(pms))
    gm = pairs * n * n / (ms * 1e-3) / 1e9
    # spot check against torch
    x = (q[:64].unsqueeze(1) ^ t[:n].unsqueeze(0)).to(torch.int32)
    lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int32, device=dev)
    d = lut[x].sum(-1)
    assert torch.equal(d.min(1).values, bd[:64]) and torch.equal(d.argmin(1).to(torch.int32), bi[:64])
    return {"gmatch_per_s": gm, "ms": ms, "popc_peak_per_s": peak.value, "frac_of_popc_peak": gm * 8e9 / peak.value,
            "n": n, "pairs": pairs}


if __name__ == "__main__":
    import json
    print(json.dumps(run(int(sys.argv[1]) if len(sys.argv) > 1 else 10000, int(sys.argv[2]) if len(sys.argv) > 2 else 32)))
