"""GPU parity of the 256-bit Hamming kernels (through the C ABI) vs the CPU oracle: pairwise DescriptorDistance and the
many-vs-many best / second-best search (src/ORBmatcher.cc:2123-2143, :325-345), up to the 10 k x 10 k size of config 4."""
import numpy as np
import pytest
import synth, oracle_lib as O
from gpu_util import pkg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    return pkg()


@pytest.fixture(scope="module")
def ctx(P):
    return P.ORBextractor(1000, 1.2, 8, 20, 7, max_w=64, max_h=64)


def popcount_rows(a, b):
    return np.unpackbits(np.bitwise_xor(a, b), axis=1).sum(axis=1).astype(np.int32)


def test_descriptor_distance_pairs(P, ctx):
    a, b = synth.descriptors(1, 5000), synth.descriptors(2, 5000)
    a[0] = 0; b[0] = 255            # 256
    a[1] = b[1]                     # 0
    b[2] = a[2]; b[2, 31] ^= 0x80   # 1
    got = P.descriptor_distance(ctx, a, b)
    assert np.array_equal(got, popcount_rows(a, b))
    assert got[0] == 256 and got[1] == 0 and got[2] == 1
    L = O._mlib()
    assert all(got[i] == L.orc_descriptor_distance(a[i].ctypes.data, b[i].ctypes.data) for i in range(0, 5000, 37))


@pytest.mark.parametrize("nq,nt,seed", [(1, 1, 0), (7, 300, 1), (1000, 33, 2), (2049, 4097, 3), (10000, 10000, 4)])
def test_best2_whole_set(P, ctx, nq, nt, seed):
    q = synth.descriptors(seed, nq)
    t = synth.descriptors(seed + 100, nt)
    m = min(nq, nt) // 2                       # true matches, some exact duplicates (ties: first train wins)
    if m:
        t[:m] = synth.descriptors(seed + 200, m, 0.05, q[:m])
        t[m // 2] = t[0]
    want = O.hamming_best2(q, t)
    got = P.hamming_best2(ctx, q, t)
    for w, g, name in zip(want, got, ("best_idx", "best_dist", "second_dist")):
        assert np.array_equal(w, g), (name, int((w != g).sum()))


def test_best2_segments_lists_and_mask(P, ctx):
    rng = np.random.default_rng(5)
    nq, nt, nseg = 3000, 3500, 40
    q, t = synth.descriptors(6, nq), synth.descriptors(7, nt)
    t[:1500] = synth.descriptors(8, 1500, 0.08, q[:1500])
    # index lists in shuffled order, segments of very different sizes including empty ones on either side
    q_idx, t_idx = rng.permutation(nq).astype(np.int32)[:2800], rng.permutation(nt).astype(np.int32)[:3300]
    cut = lambda n: np.concatenate([[0], np.sort(rng.integers(0, n + 1, nseg - 1)), [n]]).astype(np.int32)
    q_seg, t_seg = cut(len(q_idx)), cut(len(t_idx))
    q_seg[5] = q_seg[4]; t_seg[9] = t_seg[8]
    q_seg, t_seg = np.maximum.accumulate(q_seg), np.maximum.accumulate(t_seg)
    valid = (rng.random(nt) < 0.7).astype(np.uint8)
    for tv in (None, valid):
        want = O.hamming_best2(q, t, q_seg, q_idx, t_seg, t_idx, tv)
        got = P.hamming_best2(ctx, q, t, q_seg, q_idx, t_seg, t_idx, tv)
        for w, g, name in zip(want, got, ("best_idx", "best_dist", "second_dist")):
            assert np.array_equal(w, g), (name, tv is not None, int((w != g).sum()))
    none = (want[0] < 0)
    assert none.any() and np.all(want[1][none] == 256)    # rows with no candidate are reported as (-1, 256, 256)


def test_best2_self_match_property(P, ctx):
    """Size-independent property at the config-4 size: a set matched against a permutation of itself finds itself at 0."""
    n = 10000
    d = synth.descriptors(11, n)
    perm = np.random.default_rng(12).permutation(n)
    bi, bd, sd = P.hamming_best2(ctx, d, d[perm])
    assert np.all(bd == 0) and np.array_equal(perm[bi], np.arange(n)) and np.all(sd > 0)
