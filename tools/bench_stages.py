#!/usr/bin/env python3
"""Per-stage device times of the device-resident extraction step (no CPU baseline, no Hamming): the quick loop used while
tuning a kernel.  usage: bench_stages.py [--batch 1024] [--steps 10] [--w 640 --h 480 --nfeat 1000] [--e2e]"""
import argparse, ctypes, importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--w", type=int, default=640)
    ap.add_argument("--h", type=int, default=480)
    ap.add_argument("--nfeat", type=int, default=1000)
    ap.add_argument("--distinct", type=int, default=32)
    ap.add_argument("--e2e", action="store_true")
    ap.add_argument("--ini", type=int, default=20)
    ap.add_argument("--min", type=int, default=7)
    a = ap.parse_args()
    import torch, synth
    P = importlib.import_module("orb-slam2-noted_b200")
    dev = torch.device("cuda:0")
    frames = np.stack([synth.frame(s, a.w, a.h) for s in range(a.distinct)])
    reps = (a.batch + a.distinct - 1) // a.distinct
    host = torch.from_numpy(np.concatenate([frames] * reps)[:a.batch]).contiguous().pin_memory()
    ex = P.ORBextractor(a.nfeat, 1.2, 8, a.ini, a.min, device=0, max_w=a.w, max_h=a.h, max_batch=a.batch)
    cap = ex.max_keypoints()
    d_img = host.to(dev)
    d_kps = torch.empty((a.batch, cap, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.empty((a.batch, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(a.batch, dtype=torch.int32, device=dev)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)
    for _ in range(3):
        ex.extract_batch_device(d_img, d_kps, d_desc, d_n)
    ex.sync()
    ex.set_profiling(True); ex.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(a.steps):
        ex.extract_batch_device(d_img, d_kps, d_desc, d_n)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    st = ex.stage_stats()
    ex.set_profiling(False)
    out = {"ms_per_step": round(ms, 4), "frames_per_s": round(a.batch / ms * 1e3, 1),
           "stage_ms": {k: round(v[0] / a.steps, 4) for k, v in st.items() if v[0] > 0},
           "kp_per_frame": float(d_n.sum().item()) / a.batch,
           "env": {k: v for k, v in os.environ.items() if k.startswith("ORBX_")}}
    if a.e2e:
        L = P.lib()
        h_kps = torch.empty((a.batch, cap, 28), dtype=torch.uint8).pin_memory()
        h_desc = torch.empty((a.batch, cap, 32), dtype=torch.uint8).pin_memory()
        h_n = np.zeros(a.batch, np.int32)
        vp = ctypes.c_void_p

        def go():
            rc = L.orbx_extract_batch(ex._h, vp(host.data_ptr()), a.batch, a.w, a.h, a.w, a.w * a.h, 0, vp(h_kps.data_ptr()),
                                      vp(h_desc.data_ptr()), cap, h_n.ctypes.data_as(vp))
            assert rc == 0, rc
        for _ in range(2):
            go()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            go()
        torch.cuda.synchronize()
        out["e2e_frames_per_s"] = round(a.batch * 5 / (time.perf_counter() - t0), 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
