"""GPU: KeyFrameDatabase on the device (orbx_kfdb_*, csrc/k_kfdb.cu) against the oracle and, where
oracle/_ref travelled to this box, against the unmodified reference class: the same ordered candidate lists for every query
of scripts that add, erase, re-link and query up to 2200 keyframes."""
import os
import numpy as np
import pytest
import kfdb_scene as K
import ref_lib as R
from gpu_util import pkg
from test_oracle_kfdb import SCRIPTS, quirk_script

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    P = pkg()
    ex = P.ORBextractor(500, 1.2, 8, 20, 7, max_w=64, max_h=64, max_batch=1)
    yield ex
    ex.close()


def check(ops, nw, ctx):
    g, o = K.GpuBackend(nw, ctx), K.OracleBackend(nw)
    rg, ro = K.replay(ops, g), K.replay(ops, o)
    bad = [(a, b) for a, b in zip(rg, ro) if a != b]
    assert not bad and len(rg) == len(ro), bad[:3]
    if os.path.exists(R.SO):
        r = K.RefBackend(nw)
        assert K.replay(ops, r) == rg
        r.close()
    g.close(); o.close()
    return rg


@pytest.mark.parametrize("kw", SCRIPTS, ids=lambda k: f"seed{k['seed']}")
def test_kfdb_scripts_match_oracle_and_reference(kw, ctx):
    ops, nw = K.make_script(**kw)
    res = check(ops, nw, ctx)
    assert sum(len(x[2]) for x in res) > len(res)


def test_kfdb_quirks(ctx):
    ops, nw = quirk_script()
    res = check(ops, nw, ctx)
    assert [x[2] for x in res][2] == []


def test_kfdb_batched_relocalisation_and_limits(ctx):
    """Many relocalisation queries per call give what one call per query gives (the queries run in order on the stream);
    candidate lists longer than cap report their true length; malformed BowVectors are refused."""
    P = pkg()
    ops, nw = K.make_script(7, nkf=600, nqueries=0, spread=60.0, places=3, words_per_kf=150)
    g, o = K.GpuBackend(nw, ctx), K.OracleBackend(nw)
    K.replay(ops, g); K.replay(ops, o)
    rng = np.random.default_rng(5)
    qs = [K.bow_vector(rng, nw, int(rng.integers(1000, nw - 1000)), 150, 60.0) for _ in range(20)]
    base = next(op for op in ops if op[0] == "new")
    qs += [(base[2], base[3]), (np.zeros(0, np.int32), np.zeros(0)), (base[2][:40], base[3][:40])]
    ids = list(range(100, 100 + len(qs)))
    got = g.db.detect_reloc_batch(ids, qs, cap=64)
    want = [o.reloc(i, w, v) for i, (w, v) in zip(ids, qs)]
    assert got == want and any(len(x) > 0 for x in got) and got[-2] == []
    # cap smaller than the list: the count is the true one, the first cap ids are written
    full = o.reloc(500, base[2], base[3])
    if len(full) > 1:
        with pytest.raises(P.OrbxError, match="candidates"):
            g.db.detect_reloc_batch([500], [(base[2], base[3])], cap=1)
    with pytest.raises(P.OrbxError):
        g.db.add(9999, np.array([5, 3], np.int32), np.array([0.5, 0.5]))         # not ascending
    with pytest.raises(P.OrbxError):
        g.db.add(9999, np.array([nw], np.int32), np.array([1.0]))                # word outside the vocabulary
    with pytest.raises(P.OrbxError):
        g.db.add(0, base[2], base[3])                                            # already in the database
    n0 = len(g.db)
    g.db.erase(123456)                                                           # unknown id: ignored like the reference
    assert len(g.db) == n0
    g.db.clear()
    assert len(g.db) == 0 and g.db.DetectRelocalizationCandidates(900, base[2], base[3]) == []
    g.db.add(3, base[2], base[3])
    assert g.db.DetectRelocalizationCandidates(901, base[2], base[3]) == [3]
    g.close(); o.close()


def test_kfdb_thousands_of_candidates_and_degenerate_inputs(ctx):
    """3000 near-identical keyframes: every one passes the 0.8 gate, so the one-CTA final stage walks thousands of candidates
    (groups, duplicates of best keyframes, output ranks) instead of a handful; plus empty BowVectors, an empty database and
    query words that no keyframe holds."""
    P = pkg()
    nw = 5000
    rng = np.random.default_rng(11)
    base_w = np.sort(rng.choice(nw, 120, replace=False)).astype(np.int32)
    g, o = K.GpuBackend(nw, ctx), K.OracleBackend(nw)
    assert g.reloc(1, base_w, np.full(len(base_w), 1.0 / len(base_w))) == []           # empty database
    kfs = []
    for k in range(3000):
        w = base_w.copy()
        if k % 7 == 3:                                                                # a few differ in some words
            w = np.unique(np.concatenate([w[: 100 + k % 15], rng.integers(0, nw, 10).astype(np.int32)])).astype(np.int32)
        v = np.abs(np.sin(np.arange(len(w)) * 0.37 + k * 0.01)) + 0.05
        kfs.append((w, v / v.sum()))
    for b in (g, o):
        for k, (w, v) in enumerate(kfs):
            b.new(k, w, v); b.add(k)
            b.covis(k, np.array([(k * 13 + j * 101) % 3000 for j in range(k % 11)], np.int32))
        b.new(5000, np.zeros(0, np.int32), np.zeros(0)); b.add(5000)                  # a keyframe without words
    qv = np.full(len(base_w), 1.0 / len(base_w))
    for qid, (w, v) in enumerate([(base_w, qv), (base_w[::2], qv[::2] * 2), (np.zeros(0, np.int32), np.zeros(0)),
                                  (np.array([nw - 1], np.int32) if (nw - 1) not in base_w else np.array([nw - 2], np.int32), np.ones(1))], 100):
        got, want = g.reloc(qid, w, v, cap=4096), o.reloc(qid, w, v, cap=4096)
        assert got == want, (qid, len(got), len(want))
    assert len(g.reloc(200, base_w, qv, cap=4096)) > 100
    lg = g.db.DetectLoopCandidates(7000, base_w, qv, np.arange(0, 3000, 3, dtype=np.int32), 0.2, cap=4096)
    for b in (o,):
        b.new(7000, base_w, qv)
    assert lg == o.loop(7000, np.arange(0, 3000, 3, dtype=np.int32), 0.2, cap=4096) and len(lg) > 50
    g.close(); o.close()
