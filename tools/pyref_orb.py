"""Independent Python restatement of ORBextractor::operator() that calls the REAL third-party
primitives (cv2 4.13 resize / FAST / GaussianBlur / fastAtan2, host libm cosf/sinf) for every
arithmetic step, with only the control flow of src/ORBextractor.cc restated.

Used ONLY by tools/gen_golden.py (in the build container, where cv2 exists) to produce
tests/golden/*.npz.  Deliberately written differently from oracle/orb_extract.c (python lists,
per-cell cv2.FAST calls exactly as the reference makes them) so that the two pin each other.
"""
import ctypes
import math
import numpy as np
import cv2

_libm = ctypes.CDLL("libm.so.6")
_libm.cosf.restype = ctypes.c_float
_libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float
_libm.sinf.argtypes = [ctypes.c_float]
f32 = np.float32

EDGE = 19
HALF = 15
PATCH = 31


def load_pattern(path):
    txt = open(path).read().split("*/", 1)[1]
    v = [int(t) for t in txt.replace("\n", "").split(",") if t.strip()]
    return np.array(v, np.int32).reshape(512, 2)


class Params:
    def __init__(self, nfeatures, scale, nlevels, ini, mn):
        self.nfeatures, self.nlevels, self.ini, self.mn = nfeatures, nlevels, ini, mn
        sf = f32(scale)
        self.scale = [f32(1.0)]
        for _ in range(1, nlevels):
            self.scale.append(f32(self.scale[-1] * sf))
        self.inv = [f32(f32(1.0) / s) for s in self.scale]
        factor = f32(f32(1.0) / sf)
        nd = f32(f32(f32(nfeatures) * f32(f32(1) - factor)) / f32(f32(1) - f32(math.pow(float(factor), float(nlevels)))))
        self.nfeat = []
        for _ in range(nlevels - 1):
            self.nfeat.append(int(np.rint(nd)))
            nd = f32(nd * factor)
        self.nfeat.append(max(nfeatures - sum(self.nfeat), 0))
        umax = [0] * (HALF + 1)
        vmax = int(math.floor(HALF * math.sqrt(2.0) / 2 + 1))
        vmin = int(math.ceil(HALF * math.sqrt(2.0) / 2))
        for v in range(vmax + 1):
            umax[v] = int(np.rint(math.sqrt(HALF * HALF - v * v)))
        v0 = 0
        for v in range(HALF, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax


class Node:
    seq = 0

    def __init__(self, ulx, uly, brx, bry):
        self.ulx, self.uly, self.brx, self.bry = ulx, uly, brx, bry
        self.keys = []
        self.no_more = False
        self.seq = Node.seq
        Node.seq += 1

    def divide(self):
        hx = int(math.ceil(float(f32(self.brx - self.ulx) / f32(2))))
        hy = int(math.ceil(float(f32(self.bry - self.uly) / f32(2))))
        mx, my = self.ulx + hx, self.uly + hy
        n1 = Node(self.ulx, self.uly, mx, my)
        n2 = Node(mx, self.uly, self.brx, my)
        n3 = Node(self.ulx, my, mx, self.bry)
        n4 = Node(mx, my, self.brx, self.bry)
        for k in self.keys:
            if k[0] < mx:
                (n1 if k[1] < my else n3).keys.append(k)
            else:
                (n2 if k[1] < my else n4).keys.append(k)
        for n in (n1, n2, n3, n4):
            if len(n.keys) == 1:
                n.no_more = True
        return n1, n2, n3, n4


def distribute(keys, minx, maxx, miny, maxy, N):
    """keys: list of (x, y, response) window-relative, in detection order."""
    n_ini = int(np.round(f32(maxx - minx) / f32(maxy - miny)))  # std::round: no ties occur for ints ratio? handled below
    r = float(f32(maxx - minx) / f32(maxy - miny))
    n_ini = int(math.floor(r + 0.5)) if r >= 0 else -int(math.floor(-r + 0.5))
    hx = f32(f32(maxx - minx) / f32(n_ini))
    nodes = []
    for i in range(n_ini):
        nodes.append(Node(int(f32(hx * f32(i))), 0, int(f32(hx * f32(i + 1))), maxy - miny))
    ini = list(nodes)
    for k in keys:
        ini[int(f32(k[0]) / hx)].keys.append(k)
    lst = []
    for n in nodes:
        if len(n.keys) == 1:
            n.no_more = True
            lst.append(n)
        elif len(n.keys) > 1:
            lst.append(n)
    finish = False
    while not finish:
        prev = len(lst)
        expand = []
        front = []
        keep = []
        for n in lst:
            if n.no_more:
                keep.append(n)
                continue
            for c in n.divide():
                if c.keys:
                    front.insert(0, c)
                    if len(c.keys) > 1:
                        expand.append(c)
        lst = front + keep
        if len(lst) >= N or len(lst) == prev:
            finish = True
        elif len(lst) + 3 * len(expand) > N:
            while not finish:
                prev = len(lst)
                pv = sorted(expand, key=lambda c: (len(c.keys), c.seq))
                expand = []
                for n in reversed(pv):
                    for c in n.divide():
                        if c.keys:
                            lst.insert(0, c)
                            if len(c.keys) > 1:
                                expand.append(c)
                    lst.remove(n)
                    if len(lst) >= N:
                        break
                if len(lst) >= N or len(lst) == prev:
                    finish = True
    out = []
    for n in lst:
        best = n.keys[0]
        for k in n.keys[1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return out


def detect_level(img, p):
    h, w = img.shape
    minb = EDGE - 3
    maxbx, maxby = w - EDGE + 3, h - EDGE + 3
    width, height = f32(maxbx - minb), f32(maxby - minb)
    ncols, nrows = int(width / f32(30)), int(height / f32(30))
    if ncols < 1 or nrows < 1:
        return []
    wcell = int(math.ceil(float(width / f32(ncols))))
    hcell = int(math.ceil(float(height / f32(nrows))))
    det_ini = cv2.FastFeatureDetector_create(p.ini, True)
    det_min = cv2.FastFeatureDetector_create(p.mn, True)
    out = []
    for i in range(nrows):
        iniy = minb + i * hcell
        maxy = iniy + hcell + 6
        if iniy >= maxby - 3:
            continue
        maxy = min(maxy, maxby)
        for j in range(ncols):
            inix = minb + j * wcell
            maxx = inix + wcell + 6
            if inix >= maxbx - 6:
                continue
            maxx = min(maxx, maxbx)
            roi = np.ascontiguousarray(img[iniy:maxy, inix:maxx])
            kps = det_ini.detect(roi)
            if len(kps) == 0:
                kps = det_min.detect(roi)
            for k in kps:
                out.append((float(f32(k.pt[0]) + f32(j * wcell)), float(f32(k.pt[1]) + f32(i * hcell)), float(k.response)))
    return out


def ic_angle(img, cx, cy, umax):
    m01 = m10 = 0
    for v in range(-HALF, HALF + 1):
        d = umax[abs(v)]
        row = img[cy + v, cx - d:cx + d + 1].astype(np.int64)
        u = np.arange(-d, d + 1)
        m10 += int((u * row).sum())
        m01 += v * int(row.sum())
    return f32(cv2.fastAtan2(float(m01), float(m10)))


def brief(blur, cx, cy, angle, pattern):
    ang = f32(f32(angle) * f32(math.pi / float(f32(180.0))))
    a = f32(_libm.cosf(ctypes.c_float(float(ang))))
    b = f32(_libm.sinf(ctypes.c_float(float(ang))))
    px = pattern[:, 0].astype(np.float32)
    py = pattern[:, 1].astype(np.float32)
    rr = np.rint((px * b).astype(np.float32) + (py * a).astype(np.float32)).astype(np.int64)
    cc = np.rint((px * a).astype(np.float32) - (py * b).astype(np.float32)).astype(np.int64)
    vals = blur[cy + rr, cx + cc].astype(np.int32)
    bits = (vals[0::2] < vals[1::2]).astype(np.uint8)
    return np.packbits(bits, bitorder="little")


def extract(img, p, pattern, ksize=9):
    """Returns (kps float32 [n,5]: x,y,size,angle,response; octave int32 [n]; desc uint8 [n,32]; pyramid list)."""
    h0, w0 = img.shape
    pyr = []
    for l in range(p.nlevels):
        w = int(np.rint(f32(f32(w0) * p.inv[l])))
        h = int(np.rint(f32(f32(h0) * p.inv[l])))
        if l == 0:
            pyr.append(img.copy())
        else:
            pyr.append(cv2.resize(pyr[l - 1], (w, h), interpolation=cv2.INTER_LINEAR))
    rows, octs, descs = [], [], []
    cands = []
    for l in range(p.nlevels):
        im = pyr[l]
        h, w = im.shape
        keys = detect_level(im, p)
        cands.append(keys)
        if not keys:
            continue
        sel = distribute(keys, EDGE - 3, w - EDGE + 3, EDGE - 3, h - EDGE + 3, p.nfeat[l])
        if not sel:
            continue
        blur = cv2.GaussianBlur(im, (ksize, ksize), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        size = float(int(f32(PATCH) * p.scale[l]))
        for (x, y, r) in sel:
            x = f32(f32(x) + f32(EDGE - 3))
            y = f32(f32(y) + f32(EDGE - 3))
            cx, cy = int(np.rint(x)), int(np.rint(y))
            ang = ic_angle(im, cx, cy, p.umax)
            descs.append(brief(blur, cx, cy, ang, pattern))
            if l != 0:
                x = f32(x * p.scale[l])
                y = f32(y * p.scale[l])
            rows.append((x, y, size, ang, r))
            octs.append(l)
    kp = np.array(rows, np.float32).reshape(-1, 5)
    return kp, np.array(octs, np.int32), (np.stack(descs) if descs else np.zeros((0, 32), np.uint8)), pyr, cands
