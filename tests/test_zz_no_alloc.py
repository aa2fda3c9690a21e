"""GPU: orbx_hamming_best2_static -- the no-allocation variant of orbx_hamming_best2 (include/orbx.h; no counterpart in
the reference) -- called through the C ABI and compared with the existing
entry point.  Integer outputs, so the tolerance is the one every Hamming test here uses: exact equality.

Collected last on purpose (the file name sorts after every other test file): a
failure here must not hide the parity tests.  It passes alone in a fresh process and after the other tests: it creates
its own context and depends on no state but the library's static scratch array, which it deliberately dirties."""
import ctypes
import numpy as np
import pytest
import synth
from gpu_util import pkg

pytestmark = pytest.mark.gpu
IN_DEVICE, OUT_DEVICE = 1, 2
E_CAPACITY = -4
STATIC_BYTES, STATIC_MAX_SEG = 1 << 20, 1024          # ORBX_HAMMING_STATIC_BYTES / _MAX_SEG of include/orbx.h


@pytest.fixture(scope="module")
def P():
    return pkg()


@pytest.fixture(scope="module")
def ctx(P):
    return P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _call(fn, ctx, q, t, q_seg, q_idx, t_seg, t_idx, t_valid):
    nrow = len(q) if q_idx is None else len(q_idx)
    out = [np.full(nrow, -77, np.int32) for _ in range(3)]
    rc = fn(ctx._h, _p(q), len(q), _p(t), len(t), _p(q_seg), _p(q_idx), _p(t_seg), _p(t_idx), len(q_seg) - 1, _p(t_valid), 0,
            *[_p(o) for o in out])
    return rc, out


def _scene(seed, nq, nt, nseg):
    """Segments of very different sizes (some empty), shuffled index lists, a validity mask, true matches and duplicates."""
    rng = np.random.default_rng(seed)
    q, t = synth.descriptors(seed + 1, nq), synth.descriptors(seed + 2, nt)
    m = min(nq, nt) // 2
    t[:m] = synth.descriptors(seed + 3, m, 0.08, q[:m])
    t[m // 2] = t[0]
    q_idx = rng.permutation(nq).astype(np.int32)[:nq - nq // 10]
    t_idx = rng.permutation(nt).astype(np.int32)[:nt - nt // 12]
    cut = lambda n: np.concatenate([[0], np.sort(rng.integers(0, n + 1, nseg - 1)), [n]]).astype(np.int32)
    q_seg, t_seg = cut(len(q_idx)), cut(len(t_idx))
    if nseg > 9:
        q_seg[5] = q_seg[4]; t_seg[9] = t_seg[8]
        q_seg, t_seg = np.maximum.accumulate(q_seg), np.maximum.accumulate(t_seg)
    valid = (rng.random(nt) < 0.7).astype(np.uint8)
    return q, t, q_seg, q_idx, t_seg, t_idx, valid


def test_static_variant_equals_existing_entry_point_two_calls_in_a_row(P, ctx):
    L = P.lib()
    # The second scene is smaller than the first, so the shared array still holds the first call's segments,
    # descriptors and results when it runs; the third goes back up.  Each is checked against orbx_hamming_best2.
    for k, (seed, nq, nt, nseg) in enumerate([(31, 3000, 3500, 40), (32, 700, 500, 12), (33, 4000, 4100, 1)]):
        q, t, q_seg, q_idx, t_seg, t_idx, valid = _scene(seed, nq, nt, nseg)
        for tv in (valid, None):
            rc, want = _call(L.orbx_hamming_best2, ctx, q, t, q_seg, q_idx, t_seg, t_idx, tv)
            assert rc == 0, rc
            rc, got = _call(L.orbx_hamming_best2_static, ctx, q, t, q_seg, q_idx, t_seg, t_idx, tv)
            assert rc == 0, (rc, L.orbx_last_cuda_error(ctx._h))
            for w, g, name in zip(want, got, ("best_idx", "best_dist", "second_dist")):
                assert np.array_equal(w, g), (k, name, tv is not None, int((w != g).sum()))
            assert not np.any(got[1] == -77) and (got[1] < 256).any()           # every row written, real matches found


def test_static_variant_device_buffers_config4_size_and_no_allocation(P, ctx):
    """10 k x 10 k descriptors resident in HBM (beyond what the static array could stage from the host): only the three CSR
    arrays use the array.  Device memory in use must not grow over the calls (the 1 MiB array is part of the module)."""
    import torch
    L = P.lib()
    dev = torch.device("cuda", 0)
    n = 10000
    q = synth.descriptors(41, n)
    t = synth.descriptors(42, n)
    t[:5000] = synth.descriptors(43, 5000, 0.05, q[:5000])
    seg = np.array([0, 2500, 2500, 7000, n], np.int32)
    dq, dt = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
    vp = ctypes.c_void_p
    outs = {name: [torch.full((n,), -77, dtype=torch.int32, device=dev) for _ in range(3)] for name in ("want", "got")}
    torch.cuda.synchronize()

    def run(fn, o):
        return fn(ctx._h, vp(dq.data_ptr()), n, vp(dt.data_ptr()), n, _p(seg), None, _p(seg), None, len(seg) - 1, None,
                  IN_DEVICE | OUT_DEVICE, *[vp(x.data_ptr()) for x in o])
    assert run(L.orbx_hamming_best2, outs["want"]) == 0
    assert run(L.orbx_hamming_best2_static, outs["got"]) == 0          # first call: symbol address looked up (set-up)
    ctx.sync()
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(2):                                                  # two calls in a row
        for x in outs["got"]:
            x.fill_(-77)
        assert run(L.orbx_hamming_best2_static, outs["got"]) == 0
        ctx.sync()
        for w, g, name in zip(outs["want"], outs["got"], ("best_idx", "best_dist", "second_dist")):
            assert torch.equal(w, g), name
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free0 - free1 < (1 << 20), (free0, free1)                    # any cudaMalloc would take at least 1 MiB
    assert int((outs["got"][1] == 0).sum()) == 0 and int((outs["got"][1] < 60).sum()) >= 4000


def test_static_variant_rejects_sizes_beyond_the_maximum(P, ctx):
    L = P.lib()
    # (1) host staging larger than the array: 20 000 + 20 000 descriptors = 1.28 MB > ORBX_HAMMING_STATIC_BYTES
    n = 20000
    q, t = synth.descriptors(51, n), synth.descriptors(52, n)
    seg = np.array([0, n], np.int32)
    assert 2 * n * 32 > STATIC_BYTES
    rc, out = _call(L.orbx_hamming_best2_static, ctx, q, t, seg, None, seg, None, None)
    assert rc == E_CAPACITY and b"ORBX_HAMMING_STATIC_BYTES" in L.orbx_last_cuda_error(ctx._h)
    assert all(np.all(o == -77) for o in out)                           # nothing was written
    # (2) more segments than ORBX_HAMMING_STATIC_MAX_SEG
    nseg = STATIC_MAX_SEG + 1
    q, t = synth.descriptors(53, nseg), synth.descriptors(54, nseg)
    seg = np.arange(nseg + 1, dtype=np.int32)
    rc, _ = _call(L.orbx_hamming_best2_static, ctx, q, t, seg, None, seg, None, None)
    assert rc == E_CAPACITY and b"ORBX_HAMMING_STATIC_MAX_SEG" in L.orbx_last_cuda_error(ctx._h)
    # (3) exactly at the segment limit it runs, and the existing entry point agrees
    nseg = STATIC_MAX_SEG
    seg = np.arange(nseg + 1, dtype=np.int32)
    rc, want = _call(L.orbx_hamming_best2, ctx, q[:nseg], t[:nseg], seg, None, seg, None, None)
    assert rc == 0
    rc, got = _call(L.orbx_hamming_best2_static, ctx, q[:nseg], t[:nseg], seg, None, seg, None, None)
    assert rc == 0 and all(np.array_equal(w, g) for w, g in zip(want, got))
