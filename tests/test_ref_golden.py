"""Golden outputs of the UNMODIFIED reference (tests/golden/ref_outputs.json, written by tools/gen_golden_ref.py from oracle/_ref):
the oracle must reproduce them on any machine (CPU), and so must the CUDA path (GPU), whether or not the reference library
itself is present there."""
import hashlib, json, os
import numpy as np
import pytest
import synth, oracle_lib as O
import kfdb_scene as K

G = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_outputs.json")))


def digest(kps, desc):
    return hashlib.sha256(np.ascontiguousarray(kps).tobytes() + np.ascontiguousarray(desc).tobytes()).hexdigest()


def kfdb_scripts():
    from test_oracle_kfdb import SCRIPTS, quirk_script
    s = {f"seed{kw['seed']}": K.make_script(**kw) for kw in SCRIPTS if kw.get("nkf", 300) <= 400}
    s["quirks"] = quirk_script()
    return s


@pytest.mark.parametrize("name", sorted(G["extract"]))
def test_oracle_reproduces_reference_extraction(name):
    c = G["extract"][name]
    kps, desc = O.extract(O.params(c["nfeatures"]), synth.frame(c["seed"], c["w"], c["h"], c["kind"]))
    assert len(kps) == c["n"] and digest(kps, desc) == c["sha256"]


def test_oracle_reproduces_reference_keyframe_database():
    for name, (ops, nw) in kfdb_scripts().items():
        o = K.OracleBackend(nw)
        got = [[k, int(i), [int(x) for x in c]] for k, i, c in K.replay(ops, o)]
        o.close()
        assert got == G["kfdb"][name], name


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(G["extract"]))
def test_gpu_reproduces_reference_extraction(name):
    from gpu_util import pkg
    P = pkg()
    c = G["extract"][name]
    ex = P.ORBextractor(c["nfeatures"], 1.2, 8, 20, 7, max_w=c["w"], max_h=c["h"])
    kps, desc = ex(synth.frame(c["seed"], c["w"], c["h"], c["kind"]))
    ex.close()
    assert len(kps) == c["n"] and digest(kps, desc) == c["sha256"]


@pytest.mark.gpu
def test_gpu_reproduces_reference_keyframe_database():
    from gpu_util import pkg
    P = pkg()
    ex = P.ORBextractor(500, 1.2, 8, 20, 7, max_w=64, max_h=64)
    for name, (ops, nw) in kfdb_scripts().items():
        g = K.GpuBackend(nw, ex)
        got = [[k, int(i), [int(x) for x in c]] for k, i, c in K.replay(ops, g)]
        g.close()
        assert got == G["kfdb"][name], name
    ex.close()
