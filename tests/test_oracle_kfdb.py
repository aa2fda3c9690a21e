"""CPU: the KeyFrameDatabase restatement (oracle/orb_kfdb.c) against the UNMODIFIED reference class (oracle/_ref:
src/KeyFrameDatabase.cc with DBoW2's BowVector / L1Scoring) on scripts of add / erase / covisibility updates / loop and
relocalisation queries, plus hand-made cases for the reference's quirks."""
import os
import numpy as np
import pytest
import kfdb_scene as K
import ref_lib as R

needs_ref = pytest.mark.skipif(not os.path.exists(R.SO), reason="oracle/_ref not built (no reference checkout on this machine)")

SCRIPTS = [dict(seed=1), dict(seed=2, spread=120.0, places=4), dict(seed=3, spread=40.0, places=3, words_per_kf=120),
           dict(seed=4, nkf=2200, nqueries=120, spread=150.0, places=20, nwords=100000, words_per_kf=400)]


@needs_ref
@pytest.mark.parametrize("kw", SCRIPTS, ids=lambda k: f"seed{k['seed']}")
def test_oracle_matches_reference_keyframe_database(kw):
    ops, nw = K.make_script(**kw)
    o, r = K.OracleBackend(nw), K.RefBackend(nw)
    ro, rr = K.replay(ops, o), K.replay(ops, r)
    assert len(ro) >= 80 and ro == rr
    assert sum(len(x[2]) for x in ro) > len(ro)          # the scripts do find candidates
    o.close(); r.close()


def quirk_script():
    """Small hand-made database that makes the reference's corner rules visible in the output."""
    ops = []
    bow = lambda ws: (np.array(sorted(ws), np.int32), np.full(len(ws), 1.0 / len(ws)))
    kfs = {0: range(0, 40), 1: range(0, 36), 2: range(10, 50), 3: range(100, 140), 4: list(range(0, 30)) + list(range(100, 110)),
           5: range(20, 60), 6: range(0, 40)}
    for k, ws in kfs.items():
        ops += [("new", k, *bow(ws)), ("add", k)]
    ops += [("covis", 0, np.array([1, 2, 4], np.int32)), ("covis", 1, np.array([0], np.int32)), ("covis", 2, np.array([5, 0], np.int32)),
            ("covis", 6, np.array([0, 1], np.int32)), ("covis", 4, np.array([3], np.int32))]
    q = lambda ws: bow(ws)
    ops += [("reloc", 11, *q(range(100, 140))),          # scores keyframes 3 (and 4 if above the 0.8 gate)
            ("reloc", 12, *q(range(0, 40))),             # 4 shares words but stays below the gate: its score of query 11 is added to 0's group
            ("reloc", 12, *q(range(0, 40))),             # the same frame id again: nothing is listed (mnRelocQuery == id)
            ("reloc", 13, *q(range(0, 40)))]
    ops += [("new", 7, *bow(range(0, 40))),
            ("loop", 7, np.array([0, 6], np.int32), 0.0),   # connected keyframes are left out and do not feed groups
            ("loop", 7, np.array([], np.int32), 0.5),       # same query id: keyframes listed before are not listed again, the others are
            ("erase", 1), ("add", 7), ("covis", 7, np.array([0, 1, 6], np.int32)),
            ("new", 8, *bow(range(5, 45))), ("loop", 8, np.array([7], np.int32), 0.05),
            ("reloc", 14, *q(range(1000, 1010)))]        # no shared word at all
    return ops, 2000


@needs_ref
def test_quirks_oracle_matches_reference():
    ops, nw = quirk_script()
    o, r = K.OracleBackend(nw), K.RefBackend(nw)
    ro, rr = K.replay(ops, o), K.replay(ops, r)
    assert ro == rr
    by = [x[2] for x in ro]
    assert by[2] == [] and by[-1] == [] and len(by[1]) >= 1 and len(by[3]) >= 1
    o.close(); r.close()
