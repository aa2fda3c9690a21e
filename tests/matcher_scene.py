"""One two-keyframe scene for every ORBmatcher entry point, shared by the drop-in test (tests/test_gpu_cpp.py: the C++
class over the CUDA library) and the reference pin (tests/test_oracle_vs_ref.py: the UNMODIFIED src/ORBmatcher.cc over the
reference's own Frame / KeyFrame / MapPoint).  Both run the same scene file and must produce the same records."""
import numpy as np
import synth, oracle_lib as O


def _arr(f, a):
    b = np.ascontiguousarray(a).tobytes()
    f.write(np.int32(len(b)).tobytes()); f.write(b)


def _read_all(path):
    b = open(path, "rb").read()
    out, o = [], 0
    while o < len(b):
        n = int(np.frombuffer(b, np.int32, 1, o)[0]); o += 4
        out.append(b[o:o + n]); o += n
    return out


def make_pair(seed, n1, n2, flip=0.08, dup=0.0, nlevels=8, KP=None):
    """Two feature sets with true correspondences: set 2 = perturbed, shuffled subset of set 1 + random rest."""
    rng = np.random.default_rng(seed)
    d1 = synth.descriptors(seed, n1)
    if dup > 0:  # many identical descriptors: forces claimed-entry handling and node rescans
        src = rng.integers(0, max(1, int(n1 * (1 - dup))), n1)
        d1 = d1[src]
    m = min(n1, n2) * 3 // 4
    pick = rng.permutation(n1)[:m]
    d2 = np.concatenate([synth.descriptors(seed + 1, m, flip, d1[pick]), synth.descriptors(seed + 2, n2 - m)])
    k1 = synth.keypoints(seed + 3, n1, dtype=KP)
    k2 = synth.keypoints(seed + 4, n2, dtype=KP)
    rot = np.float32(37.5)
    k2["angle"][:m] = np.mod(k1["angle"][pick] - rot + rng.normal(0, 3, m).astype(np.float32), np.float32(360)).astype(np.float32)
    k2["x"][:m] = k1["x"][pick] + rng.normal(0, 2, m).astype(np.float32)
    k2["y"][:m] = k1["y"][pick] + rng.normal(0, 2, m).astype(np.float32)
    k2["octave"][:m] = k1["octave"][pick]
    perm = rng.permutation(n2)
    return d1, k1, d2[perm], k2[perm]


def proj_scene(seed, n, nq):
    rng = np.random.default_rng(seed)
    keys = synth.keypoints(seed, n, 640, 480, dtype=O.KP_DTYPE)
    desc = synth.descriptors(seed + 1, n)
    src = rng.integers(0, n, nq)
    pts = np.zeros(nq, O.PROJ_DTYPE)
    pts["u"] = keys["x"][src] + rng.normal(0, 3, nq).astype(np.float32)
    pts["v"] = keys["y"][src] + rng.normal(0, 3, nq).astype(np.float32)
    pts["u"][:4] = [-30, 900, 100, 320]; pts["v"][:4] = [100, 100, -50, 700]
    pts["level"] = np.clip(keys["octave"][src] + rng.integers(-1, 2, nq), 0, 7)
    pts["view_cos"] = np.where(rng.random(nq) < 0.5, 0.9995, 0.9).astype(np.float32)
    pts["angle"] = np.mod(keys["angle"][src] + 20 + rng.normal(0, 4, nq), 360).astype(np.float32)
    pts["desc"] = synth.descriptors(seed + 2, nq, 0.12, desc[src])
    uright = np.where(rng.random(n) < 0.6, keys["x"] - 8, -1).astype(np.float32)
    pts["ur"] = pts["u"] - 8 + rng.normal(0, 6, nq).astype(np.float32)
    occupied = (rng.random(n) < 0.1).astype(np.uint8)
    return keys, desc, uright, occupied, pts


def _similarity(rng, scale, angle_deg, t):
    """4x4 float32 [s R | t; 0 0 0 1] with a rotation of angle_deg about a random axis."""
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    a = np.deg2rad(angle_deg)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    Rm = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * (K @ K)
    T = np.eye(4)
    T[:3, :3] = scale * Rm; T[:3, 3] = t
    return T.astype(np.float32)


def check_scene(run_driver, tmp_path, seed=11, n1=1500, n2=1300, general_poses=False):
    """Builds one two-keyframe scene file, runs it through run_driver(in_path, out_path) and checks every output record
    against the oracle (candidate loops) plus a Python replay of the map updates.  Returns the raw output bytes.
    general_poses=True: the projection sections get poses with real rotations, scales and translations; the expectations
    below assume axis-aligned poses, so that scene is only run and returned (the caller compares two drivers' bytes)."""
    d1, k1, d2, k2 = make_pair(seed, n1, n2, 0.08, 0.2, KP=O.KP_DTYPE)
    v = synth.vocabulary(0, 10, 3)
    vo = O.VocabArrays(v["k"], v["L"], v["child_off"], v["child"], v["desc"], v["word_id"], v["weight"])
    rng = np.random.default_rng(seed)
    sides, nodes = [], []
    for d, k in ((d1, k1), (d2, k2)):
        w, wt, nd = O.bow_transform(vo, d, 2)
        nid, noff, feat = O.feature_vector(nd, wt)
        has = (rng.random(len(d)) < 0.6).astype(np.uint8)
        ur = np.where(rng.random(len(d)) < 0.5, k["x"] - 5, -1).astype(np.float32)
        sides.append((d, k, has, ur, nid, noff, feat))
        nodes.append(np.where(wt > 0, nd, -1).astype(np.int32))
    p = O.params(1000)
    scale = np.array(p.scale[:8], np.float32); sigma2 = (np.array(p.sigma2[:8], np.float32) * 9).astype(np.float32)
    F12 = (np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + np.float32(1e-4) * rng.normal(size=(3, 3)).astype(np.float32))
    fx, fy, cx, cy = np.float32(520.9), np.float32(521.0), np.float32(325.1), np.float32(249.7)
    Ow = np.array([0.3, -0.1, 2.0], np.float32)
    invz = np.float32(1.0) / Ow[2]
    ex = np.float32(np.float32(np.float32(fx * Ow[0]) * invz) + cx); ey = np.float32(np.float32(np.float32(fy * Ow[1]) * invz) + cy)
    nnratio, th = np.float32(0.7), np.float32(3.0)
    grid = (np.float32(0), np.float32(0), np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    # projection scene on side 2's keypoints
    keys, desc, uright, occupied, pts = proj_scene(seed + 5, n2, 900)
    keys = k2.copy(); desc = d2
    src = rng.integers(0, n2, len(pts))
    pts["desc"] = synth.descriptors(seed + 9, len(pts), 0.12, desc[src])                  # true correspondences near (u, v)
    pts["u"] = keys["x"][src] + rng.normal(0, 3, len(pts)).astype(np.float32)
    pts["v"] = keys["y"][src] + rng.normal(0, 3, len(pts)).astype(np.float32)
    pts["level"] = np.clip(keys["octave"][src] + rng.integers(-1, 2, len(pts)), 0, 7)
    uright = sides[1][3]
    # map points of KF1 placed so that they project (identity pose) next to keypoints of side 2; float32 steps mirror
    # host/ORBmatcher.cc (which mirrors src/ORBmatcher.cc:1949-1978 and :470-512)
    f32 = np.float32
    has1 = sides[0][2]
    tgt = rng.integers(0, n2, n1)
    zw = (2 + 4 * rng.random(n1)).astype(f32)
    xw = ((k2["x"][tgt] + rng.normal(0, 3, n1).astype(f32) - cx) * zw / fx).astype(f32)
    yw = ((k2["y"][tgt] + rng.normal(0, 3, n1).astype(f32) - cy) * zw / fy).astype(f32)
    xw[:6] = [9, -9, 0, 0, 0.1, 0.2]; yw[:6] = [0, 0, 9, -9, 0.1, 0.2]; zw[4] = -1.0   # out of the image / behind the camera
    lvl_w = np.clip(k2["octave"][tgt] + rng.integers(-1, 2, n1), 0, 7)
    dist = np.sqrt(xw.astype(np.float64) ** 2 + yw.astype(np.float64) ** 2 + zw.astype(np.float64) ** 2).astype(f32)
    logsf = f32(np.log(1.2))
    maxd = (dist * f32(1.2) ** (lvl_w.astype(f32) - f32(0.5))).astype(f32)
    mind = (maxd / scale[7]).astype(f32)
    maxd[6] = dist[6] / f32(1.3); mind[7] = dist[7] * f32(1.5)                            # outside the distance range
    nrm = np.stack([xw, yw, zw], 1).astype(np.float64)
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(f32)
    nrm[8::9] *= -1                                                                       # seen from behind (Sim3 variant only)
    world = np.concatenate([np.stack([xw, yw, zw, maxd, mind], 1), nrm], 1).astype(f32)
    th_reloc, orb_dist, th_sim3, window = 10.0, 100, 10, 100
    mpdesc = synth.descriptors(seed + 21, n1, 0.1, d2[tgt])
    # the same for KF2's map points, placed next to keypoints of side 1 (SearchBySim3 searches both ways)
    has2 = sides[1][2]
    tgt2 = rng.integers(0, n1, n2)
    inv = np.full(n2, -1); inv[tgt[::-1]] = np.arange(n1)[::-1]          # some i1 with tgt[i1] == i2, where one exists
    tgt2 = np.where(inv >= 0, inv, tgt2)
    zw2 = (2 + 4 * rng.random(n2)).astype(f32)
    xw2 = ((k1["x"][tgt2] + rng.normal(0, 2, n2).astype(f32) - cx) * zw2 / fx).astype(f32)
    yw2 = ((k1["y"][tgt2] + rng.normal(0, 2, n2).astype(f32) - cy) * zw2 / fy).astype(f32)
    lvl_w2 = np.clip(k1["octave"][tgt2] + rng.integers(0, 2, n2), 0, 7)
    dist2 = np.sqrt(xw2.astype(np.float64) ** 2 + yw2.astype(np.float64) ** 2 + zw2.astype(np.float64) ** 2).astype(f32)
    maxd2 = (dist2 * f32(1.2) ** (lvl_w2.astype(f32) - f32(0.5))).astype(f32)
    mind2 = (maxd2 / scale[7]).astype(f32)
    nrm2 = np.stack([xw2, yw2, zw2], 1).astype(np.float64)
    nrm2 = (nrm2 / np.linalg.norm(nrm2, axis=1, keepdims=True)).astype(f32)
    world2 = np.concatenate([np.stack([xw2, yw2, zw2, maxd2, mind2], 1), nrm2], 1).astype(f32)
    mpdesc2 = synth.descriptors(seed + 22, n2, 0.1, d1[tgt2])
    obs_n = rng.integers(1, 6, n1 + n2).astype(np.int32)
    th_fuse, bf_fuse, th_fuse_s, th_s3 = 3.0, 40.0, 4.0, 7.5
    # (d) a third of the projected points and (k) a third of the last frame's points are temporal ones (Observations() == 0,
    # Tracking::UpdateLastFrame): their matches do not hide the keypoint from the following points (:136-138, :1840-1842)
    proj_obs = np.where(rng.random(len(pts)) < 0.35, 0, rng.integers(1, 5, len(pts))).astype(np.int32)
    last_obs = np.where(rng.random(n1) < 0.35, 0, rng.integers(1, 5, n1)).astype(np.int32)
    th_motion = 15.0
    if general_poses:
        Tck = _similarity(rng, 1.0, 2.0, [0.03, -0.02, -0.12])               # (k) current frame: small rotation, forward motion
        Scw_f = _similarity(rng, 1.7, 1.5, [0.05, 0.02, 0.04])               # (f) Sim3 projection
        Scw_i = _similarity(rng, 0.6, 1.0, [-0.03, 0.02, 0.05])              # (i) Sim3 fuse
        Tcr = _similarity(rng, 1.0, 1.5, [0.04, 0.01, -0.03])                # (e) relocalisation frame
    else:
        Tck = np.eye(4, dtype=np.float32); Tck[:3, 3] = [0.01, -0.005, -0.1]
        Scw_f = Scw_i = Tcr = np.zeros(0, np.float32)                        # empty = the drivers' axis-aligned defaults
    inp, outp = tmp_path / "m_in.bin", tmp_path / "m_out.bin"
    with open(inp, "wb") as f:
        _arr(f, np.array([nnratio, 1, 0, fx, fy, cx, cy, th, grid[2], grid[3], 0, 0, 640, 480, 0, 0], np.float32))
        for (d, k, has, ur, _, _, _), nd in zip(sides, nodes):
            _arr(f, d); _arr(f, k); _arr(f, has); _arr(f, ur); _arr(f, nd)
        _arr(f, scale); _arr(f, sigma2); _arr(f, F12); _arr(f, Ow)
        _arr(f, np.stack([pts["u"], pts["v"], pts["ur"], pts["view_cos"], pts["level"].astype(np.float32)], 1).astype(np.float32))
        _arr(f, pts["desc"]); _arr(f, occupied)
        _arr(f, world); _arr(f, np.array([th_reloc, orb_dist, th_sim3, window, logsf, 2.0, 0, 0], np.float32)); _arr(f, mpdesc)
        _arr(f, world2); _arr(f, mpdesc2); _arr(f, obs_n); _arr(f, np.array([th_fuse, bf_fuse, th_fuse_s, th_s3, 0, 0, 0, 0], np.float32))
        _arr(f, proj_obs); _arr(f, last_obs); _arr(f, Tck); _arr(f, np.array([th_motion, 0, 0, 0], np.float32))      # 25..28
        _arr(f, Scw_f); _arr(f, Scw_i); _arr(f, Tcr)                                                                 # 29..31
    run_driver(str(inp), str(outp))
    if general_poses:
        return open(outp, "rb").read()
    out = _read_all(outp)
    so = [O.Side(*s) for s in sides]
    want, nw = O.search_by_bow(so[0], so[1], float(nnratio), True, False)
    assert int(np.frombuffer(out[0], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[1], np.int32), want)
    want, nw = O.search_by_bow(so[0], so[1], float(nnratio), True, True)
    assert int(np.frombuffer(out[2], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[3], np.int32), want)
    sc16 = np.zeros(16, np.float32); sc16[:8] = scale
    sg16 = np.zeros(16, np.float32); sg16[:8] = sigma2
    want, nw = O.search_for_triangulation(so[0], so[1], F12, float(ex), float(ey), sc16, sg16, False, True)
    assert int(np.frombuffer(out[4], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[5], np.int32), want)
    assert nw > 0
    want, nw = O.search_by_projection(pts, keys, desc, uright, scale, grid, occupied, float(th), nnratio=float(nnratio),
                                      transient=(proj_obs == 0))
    assert int(np.frombuffer(out[6], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[7], np.int32), want)
    assert nw > 0
    blocking, _ = O.search_by_projection(pts, keys, desc, uright, scale, grid, occupied, float(th), nnratio=float(nnratio))
    if n1 >= 1500:
        assert not np.array_equal(blocking, want) and nw > int((want >= 0).sum())   # the rule matters here: keypoints were re-taken
    assert int(np.frombuffer(out[8], np.int32)[0]) == int(np.unpackbits(d1[0] ^ d2[0]).sum())

    # the host-side projection of both overloads, restated step by step in float32
    invz = (1.0 / zw.astype(np.float64)).astype(f32)
    u_r = ((fx * xw) * invz + cx).astype(f32); v_r = ((fy * yw) * invz + cy).astype(f32)       # :1962-1963
    lvl_pred = np.clip(np.ceil(np.log((maxd / dist).astype(f32).astype(np.float64)) / np.float64(logsf)), 0, 7).astype(np.int32)
    in_range = ~((dist < f32(0.8) * mind) | (dist > f32(1.2) * maxd))
    found = np.zeros(n1, bool); found[::17] = True
    sel = (has1 != 0) & ~found & ~((u_r < 0) | (u_r > 640)) & ~((v_r < 0) | (v_r > 480)) & in_range
    idx_r = np.flatnonzero(sel)
    pr = np.zeros(len(idx_r), O.PROJ_DTYPE)
    pr["u"], pr["v"], pr["level"], pr["angle"], pr["desc"] = u_r[idx_r], v_r[idx_r], lvl_pred[idx_r], k1["angle"][idx_r], mpdesc[idx_r]
    want, nw = O.search_by_projection_reloc(pr, keys, desc, scale, grid, occupied, th_reloc, orb_dist, True)
    want = np.where(want >= 0, idx_r[np.maximum(want, 0)], -1)
    assert int(np.frombuffer(out[9], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[10], np.int32), want)
    assert nw > min(50, n1 // 30)
    u_s = (fx * (xw * invz) + cx).astype(f32); v_s = (fy * (yw * invz) + cy).astype(f32)         # :487-491
    f64 = np.float64                                                  # cv::Mat::dot accumulates in double
    dots = xw.astype(f64) * nrm[:, 0] + yw.astype(f64) * nrm[:, 1] + zw.astype(f64) * nrm[:, 2]
    sel = (has1 != 0) & ~(zw < 0) & (u_s >= 0) & (u_s < 640) & (v_s >= 0) & (v_s < 480) & in_range & ~(dots < 0.5 * dist.astype(f64))
    idx_s = np.flatnonzero(sel)
    ps = np.zeros(len(idx_s), O.PROJ_DTYPE)
    ps["u"], ps["v"], ps["level"], ps["desc"] = u_s[idx_s], v_s[idx_s], lvl_pred[idx_s], mpdesc[idx_s]
    want, nw = O.search_by_projection_sim3(ps, keys, desc, scale, grid, occupied, th_sim3)
    want = np.where(want >= 0, idx_s[np.maximum(want, 0)], -1)
    assert int(np.frombuffer(out[11], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[12], np.int32), want)
    assert nw > min(50, n1 // 30)
    prev0 = np.stack([k1["x"], k1["y"]], 1).astype(f32)
    want, nw, wprev = O.search_for_initialization(k1, d1, prev0, keys, desc, grid, window, float(nnratio), True)
    assert int(np.frombuffer(out[13], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[14], np.int32), want)
    assert np.array_equal(np.frombuffer(out[15], f32).reshape(-1, 2).view(np.uint32), wprev.view(np.uint32))


    # ---- Fuse x2 and SearchBySim3: host-side gates restated in float32, candidate loops from the oracle, and the map
    #      updates replayed in the reference's order on a small Python model of MapPoint / KeyFrame ----
    def gates(xw_, yw_, zw_, maxd_, mind_, nrm_, key_w=640, key_h=480):
        invz_ = (1.0 / zw_.astype(np.float64)).astype(f32)
        u_ = (fx * (xw_ * invz_) + cx).astype(f32); v_ = (fy * (yw_ * invz_) + cy).astype(f32)
        dist_ = np.sqrt(xw_.astype(np.float64) ** 2 + yw_.astype(np.float64) ** 2 + zw_.astype(np.float64) ** 2).astype(f32)
        ok = ~(zw_ < 0) & (u_ >= 0) & (u_ < key_w) & (v_ >= 0) & (v_ < key_h) & ~((dist_ < f32(0.8) * mind_) | (dist_ > f32(1.2) * maxd_))
        dots_ = xw_.astype(f64) * nrm_[:, 0] + yw_.astype(f64) * nrm_[:, 1] + zw_.astype(f64) * nrm_[:, 2]
        normal_ok = ~(dots_ < 0.5 * dist_.astype(f64))
        lvl_ = np.clip(np.ceil(np.log((maxd_ / dist_).astype(f32).astype(np.float64)) / np.float64(logsf)), 0, 7).astype(np.int32)
        return dict(u=u_, v=v_, invz=invz_, ok=ok, normal_ok=normal_ok, level=lvl_)
    G1 = gates(xw, yw, zw, maxd, mind, nrm)
    G2 = gates(xw2, yw2, zw2, maxd2, mind2, nrm2)
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(f32)

    class MP:
        def __init__(self, code, nobs, kf, idx):
            self.code, self.nobs, self.bad, self.obs = code, int(nobs), False, ({kf: idx} if kf is not None else {})

    class World:
        def __init__(self):
            self.p1 = [MP(i, obs_n[i], "k1", i) if has1[i] else MP(i, 1, None, None) for i in range(n1)]
            self.p2 = [MP(100000 + j, obs_n[n1 + j], "k2", j) if has2[j] else MP(100000 + j, 1, None, None) for j in range(n2)]
            self.maps = {"k1": [self.p1[i] if has1[i] else None for i in range(n1)], "k2": [self.p2[j] if has2[j] else None for j in range(n2)]}
            self.ur = {"k1": sides[0][3], "k2": sides[1][3]}

        def add_obs(self, mp, kf, idx):                     # MapPoint::AddObservation, src/MapPoint.cc:139-150: +2 for a stereo keypoint
            if kf not in mp.obs:
                mp.obs[kf] = idx; mp.nobs += 2 if self.ur[kf][idx] >= 0 else 1

        def replace(self, this, other):                     # MapPoint::Replace, src/MapPoint.cc:257-311 (observation bookkeeping)
            if this is other:
                return
            this.bad = True
            obs, this.obs = this.obs, {}
            for kf, idx in obs.items():
                if kf not in other.obs:
                    self.maps[kf][idx] = other; self.add_obs(other, kf, idx)
                else:
                    self.maps[kf][idx] = None

    def queries(idx_list, G, mpd, ang=None, bf=None):
        pq = np.zeros(len(idx_list), O.PROJ_DTYPE)
        ii = np.array(idx_list, np.int64)
        pq["u"], pq["v"], pq["level"], pq["desc"] = G["u"][ii], G["v"][ii], G["level"][ii], mpd[ii]
        if bf is not None:
            pq["ur"] = (G["u"][ii] - f32(bf) * G["invz"][ii]).astype(f32)
        return pq

    code = lambda m: -1 if m is None else m.code
    # (h) Fuse(KF2, KF1's points, th)
    Wd = World()
    vp = [Wd.p1[i] if has1[i] else None for i in range(n1)]
    vp += [vp[i] for i in range(0, n1 - 50, 50)]
    sel = [pos for pos, m in enumerate(vp) if m is not None and G1["ok"][m.code] and G1["normal_ok"][m.code]]
    pq = queries([vp[pos].code for pos in sel], G1, mpdesc, bf=bf_fuse)
    bi, _ = O.window_best(pq, keys, desc, uright, scale, grid, th_fuse, inv_sigma2, 50)
    nf = 0
    for pos, b in zip(sel, bi):
        m = vp[pos]
        if m.bad or "k2" in m.obs or b < 0:
            continue
        inkf = Wd.maps["k2"][b]
        if inkf is not None:
            if not inkf.bad:
                if inkf.nobs > m.nobs:
                    Wd.replace(m, inkf)
                else:
                    Wd.replace(inkf, m)
        else:
            Wd.add_obs(m, "k2", int(b)); Wd.maps["k2"][b] = m
        nf += 1
    assert int(np.frombuffer(out[16], np.int32)[0]) == nf and nf > min(30, n1 // 50)
    assert list(np.frombuffer(out[17], np.int32)) == [code(m) for m in Wd.maps["k2"]]
    st = np.frombuffer(out[18], np.int32)
    assert list(st[:n1] & 1) == [int(m.bad) for m in Wd.p1] and list(st[n1:] & 1) == [int(m.bad) for m in Wd.p2]
    assert [int(x) >> 1 for x in st[:n1]] == [m.nobs for m in Wd.p1]
    assert sum(m.bad for m in Wd.p1) > 0 and sum(m.bad for m in Wd.p2) > 0        # both Replace directions happened
    # (i) Fuse(KF2, Scw, points, th, vpReplacePoint)
    Wd = World()
    vpts = [Wd.p1[i] for i in range(n1) if has1[i]]
    vpts += [vpts[i] for i in range(0, len(vpts) - 40, 40)]
    sel = [pos for pos, m in enumerate(vpts) if G1["ok"][m.code] and G1["normal_ok"][m.code]]
    pq = queries([vpts[pos].code for pos in sel], G1, mpdesc)
    bi, _ = O.window_best(pq, keys, desc, uright, scale, grid, th_fuse_s, None, 50)
    repl = [None] * len(vpts)
    nf = 0
    for pos, b in zip(sel, bi):
        if b < 0:
            continue
        inkf = Wd.maps["k2"][b]
        if inkf is not None:
            if not inkf.bad:
                repl[pos] = inkf
        else:
            Wd.add_obs(vpts[pos], "k2", int(b)); Wd.maps["k2"][b] = vpts[pos]
        nf += 1
    assert int(np.frombuffer(out[19], np.int32)[0]) == nf and nf > min(30, n1 // 50)
    assert list(np.frombuffer(out[20], np.int32)) == [code(m) for m in Wd.maps["k2"]]
    assert list(np.frombuffer(out[21], np.int32)) == [code(m) for m in repl]
    # (j) SearchBySim3(KF1, KF2, vpMatches12, 1, I, 0, th)
    m12 = [None] * n1
    done1, done2 = np.zeros(n1, bool), np.zeros(n2, bool)
    for i in range(0, n1, 23):
        if has1[i] and i < n2 and has2[i]:
            m12[i] = 100000 + i; done1[i] = True; done2[i] = True
    s1 = [i for i in range(n1) if has1[i] and not done1[i] and G1["ok"][i]]
    b1, _ = O.window_best(queries(s1, G1, mpdesc), keys, desc, uright, scale, grid, th_s3, None, 100)
    s2 = [j for j in range(n2) if has2[j] and not done2[j] and G2["ok"][j]]
    b2, _ = O.window_best(queries(s2, G2, mpdesc2), k1, d1, sides[0][3], scale, grid, th_s3, None, 100)
    vn1 = np.full(n1, -1); vn1[s1] = b1
    vn2 = np.full(n2, -1); vn2[s2] = b2
    nfound = 0
    for i1 in range(n1):
        if vn1[i1] >= 0 and vn2[vn1[i1]] == i1:
            m12[i1] = 100000 + int(vn1[i1]); nfound += 1
    assert int(np.frombuffer(out[22], np.int32)[0]) == nfound and nfound > min(20, n1 // 75)
    assert list(np.frombuffer(out[23], np.int32)) == [-1 if m is None else m for m in m12]
    # (k) SearchByProjection(CurrentFrame, LastFrame, th, bMono): the projection of src/ORBmatcher.cc:1762-1808 restated in
    #     float32 for the axis-aligned pose of this scene (Rcw = I: x3Dc = x3Dw + tcw), then the oracle's loop
    tck = Tck[:3, 3]
    outl = (np.arange(n1) % 13) == 5
    xc = (xw + tck[0]).astype(f32); yc = (yw + tck[1]).astype(f32); zc = (zw + tck[2]).astype(f32)
    invzc = (1.0 / zc.astype(np.float64)).astype(f32)
    u_k = ((fx * xc) * invzc + cx).astype(f32); v_k = ((fy * yc) * invzc + cy).astype(f32)
    sel = (has1 != 0) & ~outl & ~(invzc < 0) & ~((u_k < 0) | (u_k > 640)) & ~((v_k < 0) | (v_k > 480))
    idx_k = np.flatnonzero(sel)
    pk = np.zeros(len(idx_k), O.PROJ_DTYPE)
    pk["u"], pk["v"], pk["level"], pk["angle"], pk["desc"] = u_k[idx_k], v_k[idx_k], k1["octave"][idx_k], k1["angle"][idx_k], mpdesc[idx_k]
    pk["ur"] = (u_k[idx_k] - (f32(bf_fuse) * invzc[idx_k]).astype(f32)).astype(f32)
    assert -tck[2] > f32(bf_fuse) / fx                                 # tlc.z > mb: "forward" when not monocular (:1768)
    for r, (mono, motion) in enumerate(((False, 1), (True, 0))):
        want, nw = O.search_by_projection(pk, keys, desc, uright, scale, grid, occupied, th_motion, motion=motion, check_ori=True,
                                          transient=(last_obs[idx_k] == 0))
        want = np.where(want >= 0, idx_k[np.maximum(want, 0)], -1)
        assert int(np.frombuffer(out[24 + 2 * r], np.int32)[0]) == nw and np.array_equal(np.frombuffer(out[25 + 2 * r], np.int32), want), (mono, nw)
        assert nw > min(15, n1 // 60)
        blocking, nb = O.search_by_projection(pk, keys, desc, uright, scale, grid, occupied, th_motion, motion=motion, check_ori=True)
        if n1 >= 1500:                                                  # in the large scene keypoints do get re-taken
            assert nb != nw or not np.array_equal(np.where(blocking >= 0, idx_k[np.maximum(blocking, 0)], -1), want)
    return open(outp, "rb").read()
