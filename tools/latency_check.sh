python - <<'PY'
import ctypes, importlib, os, sys, time
import numpy as np
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "tests")]
import synth
P = importlib.import_module("orb-slam2-noted_b200")
imgs = [synth.frame(i, 640, 480) for i in range(8)]
ex = P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=640, max_h=480)
ref = [ex(im) for im in imgs]                 # eager / capture calls
for rep in range(3):
    for im, (k0, d0) in zip(imgs, ref):
        k, d = ex(im)
        assert np.array_equal(k, k0) and np.array_equal(d, d0)
t = []
for i in range(300):
    t0 = time.perf_counter(); ex(imgs[i % 8]); t.append(time.perf_counter() - t0)
print("single frame median ms:", round(float(np.median(t[50:])) * 1e3, 4), "launches", ex.kernel_launches())
PY
