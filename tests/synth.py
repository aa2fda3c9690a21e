"""Seeded synthetic inputs (numpy only, so they are identical here and on the GPU box).

The default frame kind is a scene-like image (smooth multi-octave shading + overlapping rotated rectangles + sensor
noise) whose FAST corner density (~3 % at threshold 7) resembles a textured indoor frame.  SURVEY.md 8(d)'s first
proposal (quarter-resolution noise upsampled x4) turned out to make 45 % of all pixels FAST corners; it is kept as
kind="blocks", a stress case for the dense paths.
"""
import numpy as np


def _smooth(img):
    k = np.array([1, 4, 6, 4, 1], dtype=np.uint32)
    p = np.pad(img.astype(np.uint32), 2, mode="reflect")
    h = sum(k[i] * p[:, i:i + img.shape[1]] for i in range(5))
    v = sum(k[i] * h[i:i + img.shape[0], :] for i in range(5))
    return ((v + 128) >> 8).astype(np.uint8)


def _interp_up(n, s):
    x = np.arange(n) / s
    i0 = np.floor(x).astype(np.int64)
    return i0, x - i0


def _natural(rng, w, h):
    """Scene-like frame: multi-octave smooth shading (1/f-ish), ~150 overlapping rotated rectangles per 640x480 of
    area (opaque or additive, random contrast), sensor noise sigma 1.2.  About 3 % of pixels are FAST corners at
    threshold 7 and 1 % at 20, like a textured indoor frame; float64 numpy only, so it is bit-identical everywhere."""
    img = np.full((h, w), 118.0)
    for s, amp in ((96, 38.0), (48, 26.0), (24, 17.0), (12, 10.0), (6, 5.0), (3, 2.5)):
        g = rng.standard_normal((h // s + 3, w // s + 3))
        iy, fy = _interp_up(h, s)
        ix, fx = _interp_up(w, s)
        rows = g[iy] * (1 - fy)[:, None] + g[iy + 1] * fy[:, None]
        img += amp * (rows[:, ix] * (1 - fx)[None, :] + rows[:, ix + 1] * fx[None, :])
    for _ in range(int(150 * (w * h) / (640 * 480)) + 2):
        cx, cy = rng.random() * w, rng.random() * h
        big = rng.random() < 0.35
        a = rng.random() * (90 if big else 22) + 3
        b = rng.random() * (70 if big else 18) + 3
        th = rng.random() * np.pi
        c, s_ = np.cos(th), np.sin(th)
        r = int(np.ceil(np.hypot(a, b))) + 1
        x0, x1 = max(0, int(cx) - r), min(w, int(cx) + r + 1)
        y0, y1 = max(0, int(cy) - r), min(h, int(cy) + r + 1)
        delta = (rng.random() * 70 + 25) * (1 if rng.random() < 0.5 else -1)
        mode = rng.random()
        if x1 <= x0 or y1 <= y0:
            continue
        yy, xx = np.mgrid[y0:y1, x0:x1]
        u = (xx - cx) * c + (yy - cy) * s_
        v = -(xx - cx) * s_ + (yy - cy) * c
        m = (np.abs(u) < a) & (np.abs(v) < b)
        sub = img[y0:y1, x0:x1]
        if mode < 0.5:
            sub[m] = sub[m] * 0.35 + (118 + delta) * 0.65
        else:
            sub[m] += delta * 0.6
    img += rng.standard_normal((h, w)) * 1.2
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def frame(seed, w=640, h=480, kind="texture"):
    """Return an (h, w) uint8 image.  kinds: texture (scene-like, the bench workload), blocks (4x4 blocky noise +
    rectangles: ~45 % of pixels are FAST corners, a stress case), noise, gradient, black."""
    rng = np.random.default_rng(seed)
    if kind == "black":
        return np.zeros((h, w), np.uint8)
    if kind == "texture":
        return np.ascontiguousarray(_natural(rng, w, h))
    if kind == "gradient":  # weak corners only: forces the minThFAST retry
        yy, xx = np.mgrid[0:h, 0:w]
        img = ((xx * 3 + yy * 2) // 8 % 256).astype(np.uint8)
        n = rng.integers(0, 18, size=(h // 8 + 1, w // 8 + 1)).astype(np.uint8)
        img = img + np.kron(n, np.ones((8, 8), np.uint8))[:h, :w]
        return img.astype(np.uint8)
    if kind == "noise":
        return rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    assert kind == "blocks"
    q = rng.integers(0, 256, size=((h + 3) // 4, (w + 3) // 4), dtype=np.uint8)
    img = np.kron(q, np.ones((4, 4), np.uint8))[:h, :w]
    img = _smooth(img)
    for _ in range(50):
        x0, y0 = int(rng.integers(0, w - 8)), int(rng.integers(0, h - 8))
        rw, rh = int(rng.integers(4, max(5, w // 6))), int(rng.integers(4, max(5, h // 6)))
        val = int(rng.integers(0, 256))
        if rng.random() < 0.7:
            img[y0:y0 + rh, x0:x0 + rw] = val
        else:  # thick diagonal line
            n = min(rw, rh)
            for t in range(n):
                img[min(h - 1, y0 + t):min(h, y0 + t + 3), min(w - 1, x0 + t):min(w, x0 + t + 3)] = val
    return np.ascontiguousarray(img)


def descriptors(seed, n, flip_p=None, base=None):
    """Random 256-bit descriptors (n,32) uint8; with base/flip_p: perturbed copies of base."""
    rng = np.random.default_rng(seed)
    if base is None:
        return rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    bits = np.unpackbits(base[:n], axis=1)
    flips = (rng.random(bits.shape) < flip_p).astype(np.uint8)
    return np.packbits(bits ^ flips, axis=1)


def vocabulary(seed, k=10, L=3):
    """Synthetic DBoW2 tree in flattened form (child_off, child, desc, word_id, weight), node ids in the
    ORBvoc text-file order used by the reference's own vocabularies (breadth-first creation order is not
    required by the loader; we emit depth-first like DBoW2's saver).  Children are perturbed copies of their
    parent so that descents are meaningful; a few leaves get weight 0 (stop words)."""
    rng = np.random.default_rng(seed)
    parent, desc, leaf = [-1], [np.zeros(32, np.uint8)], [0]

    def grow(pid, level, base):
        for _ in range(k):
            bits = np.unpackbits(base)
            flips = (rng.random(256) < (0.5 if level == 1 else 0.18)).astype(np.uint8)
            d = np.packbits(bits ^ flips)
            nid = len(parent)
            parent.append(pid); desc.append(d); leaf.append(1 if level == L else 0)
            if level < L:
                grow(nid, level + 1, d)
    grow(0, 1, rng.integers(0, 256, 32, dtype=np.uint8))
    n = len(parent)
    parent = np.array(parent)
    off = np.zeros(n + 1, np.int32)
    for i in range(1, n):
        off[parent[i] + 1] += 1
    off = np.cumsum(off).astype(np.int32)
    fill = off[:-1].copy()
    child = np.zeros(n - 1, np.int32)
    for i in range(1, n):
        child[fill[parent[i]]] = i
        fill[parent[i]] += 1
    word = np.zeros(n, np.int32)
    weight = np.zeros(n, np.float32)
    w = 0
    for i in range(1, n):
        if leaf[i]:
            word[i] = w
            w += 1
            weight[i] = 0.0 if rng.random() < 0.02 else np.float32(rng.random() * 3 + 0.01)
    return dict(k=k, L=L, child_off=off, child=child, desc=np.stack(desc), word_id=word, weight=weight, parent=parent, leaf=np.array(leaf))


def vocabulary_text(v):
    """ORBvoc text format (TemplatedVocabulary.h:1385-1487)."""
    lines = [f"{v['k']} {v['L']} 0 0"]
    for i in range(1, len(v["parent"])):
        lines.append(f"{v['parent'][i]} {v['leaf'][i]} " + " ".join(str(int(b)) for b in v["desc"][i]) + f" {float(v['weight'][i])!r}")
    return "\n".join(lines) + "\n"


def keypoints(seed, n, w=640, h=480, nlevels=8, dtype=None):
    """Random undistorted keypoints (x, y, angle, octave) for matcher tests."""
    rng = np.random.default_rng(seed)
    k = np.zeros(n, dtype)
    k["x"] = (rng.random(n) * (w - 40) + 20).astype(np.float32)
    k["y"] = (rng.random(n) * (h - 40) + 20).astype(np.float32)
    k["angle"] = (rng.random(n) * 360).astype(np.float32)
    k["octave"] = rng.integers(0, nlevels, n)
    k["size"] = 31
    k["class_id"] = -1
    return k


def stereo_pair(seed, w, h, max_disp=40):
    """Rectified synthetic pair: the right image is the left one shifted by a disparity that is
    constant inside 40-row bands (so corners survive), plus +-2 grey levels of noise."""
    rng = np.random.default_rng(seed + 7919)
    base = frame(seed, w + 2 * max_disp + 8, h)
    left = np.ascontiguousarray(base[:, max_disp:max_disp + w])
    right = np.empty_like(left)
    for y0 in range(0, h, 40):
        d = int(rng.integers(3, max_disp))
        right[y0:y0 + 40] = base[y0:y0 + 40, max_disp + d:max_disp + d + w]
    noise = rng.integers(-2, 3, size=right.shape)
    right = np.clip(right.astype(np.int32) + noise, 0, 255).astype(np.uint8)
    return left, np.ascontiguousarray(right)


def vocabulary_bfs(seed, k=10, L=6):
    """A full k-ary DBoW2 tree of depth L in breadth-first node order (node i's children are k*i+1 .. k*i+k), built with
    vectorised numpy so that the ORBvoc-sized k=10, L=6 tree (1 111 111 nodes, 10^6 words) takes a second.  Children are
    perturbed copies of their parent (half the bits at level 1, one bit in eight below); leaf weights are multiples of
    1/1024 (exact in float32, double and a 10-digit decimal text), 2 % of them 0 (stop words).  Same dict as vocabulary()."""
    rng = np.random.default_rng(seed)
    levels = [rng.integers(0, 256, (1, 32), dtype=np.uint8)]
    for l in range(1, L + 1):
        par = np.repeat(levels[-1], k, axis=0)
        n = len(par)
        flips = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        if l > 1:
            flips &= rng.integers(0, 256, (n, 32), dtype=np.uint8)
            flips &= rng.integers(0, 256, (n, 32), dtype=np.uint8)
        levels.append(par ^ flips)
    levels[0] = np.zeros((1, 32), np.uint8)            # the root carries no descriptor
    desc = np.concatenate(levels)
    n = len(desc)
    first_leaf = n - len(levels[-1])
    idx = np.arange(n, dtype=np.int64)
    parent = np.where(idx > 0, (idx - 1) // k, -1)
    leaf = (idx >= first_leaf).astype(np.int64)
    child_off = np.minimum(idx * k, n - 1)             # BFS: children of i start at k*i+1, i.e. position k*i in child[]
    child_off = np.concatenate([child_off, [n - 1]]).astype(np.int32)
    child = np.arange(1, n, dtype=np.int32)
    word = np.where(leaf > 0, idx - first_leaf, 0).astype(np.int32)
    wq = rng.integers(10, 3083, n)
    wq[rng.random(n) < 0.02] = 0
    weight = np.where(leaf > 0, wq / 1024.0, 0.0).astype(np.float32)
    return dict(k=k, L=L, child_off=child_off, child=child, desc=desc, word_id=word, weight=weight, parent=parent, leaf=leaf,
                weight_q=np.where(leaf > 0, wq, 0))


def vocabulary_text_fixed(v):
    """ORBvoc text format (TemplatedVocabulary.h:1385-1487) as bytes, written with fixed-width fields from numpy arrays
    (the loader reads whitespace-separated numbers, so padding is free).  Needs v["weight_q"] (weights in 1/1024)."""
    n = len(v["parent"]) - 1
    par = v["parent"][1:].astype(np.int64)
    W = 8 + 2 + 32 * 4 + 1 + 13 + 1
    buf = np.full((n, W), ord(" "), np.uint8)
    for j in range(8):
        buf[:, 7 - j] = ord("0") + (par // 10 ** j) % 10
    buf[:, 9] = ord("0") + v["leaf"][1:]
    d = v["desc"][1:].astype(np.int64)
    for j in range(3):
        buf[:, 10 + 3 - j:10 + 128:4] = ord("0") + (d // 10 ** j) % 10
    q = v["weight_q"][1:].astype(np.int64)             # q / 1024 = q * 9765625 / 10^10
    ip, fr = q // 1024, (q % 1024) * 9765625
    o = 10 + 128 + 1
    buf[:, o] = ord("0") + ip
    buf[:, o + 1] = ord(".")
    for j in range(10):
        buf[:, o + 11 - j] = ord("0") + (fr // 10 ** j) % 10
    buf[:, W - 1] = ord("\n")
    return f"{v['k']} {v['L']} 0 0\n".encode() + buf.tobytes()
