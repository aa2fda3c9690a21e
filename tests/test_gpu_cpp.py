"""GPU: the C++ drop-in class ORB_SLAM2::ORBextractor (include/ORBextractor.h, host/ORBextractor.cc) called the way
Frame::ExtractORB calls it, against the oracle: keypoints, descriptors and the bordered mvImagePyramid."""
import os, subprocess
import numpy as np
import pytest
import synth, oracle_lib as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_dropin_extractor(tmp_path):
    subprocess.run(["make", "-C", os.path.join(ROOT, "host"), "-s"], check=True)
    w, h, nf = 752, 480, 1200
    img = synth.frame(12, w, h)
    raw, out = tmp_path / "in.raw", tmp_path / "out.bin"
    img.tofile(raw)
    r = subprocess.run([os.path.join(ROOT, "host", "orbx_cpp_driver"), str(raw), str(w), str(h), str(nf), str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    b = open(out, "rb").read()
    n = int(np.frombuffer(b, np.int32, 1, 0)[0])
    o = 4
    kps = np.frombuffer(b, O.KP_DTYPE, n, o); o += 28 * n
    desc = np.frombuffer(b, np.uint8, 32 * n, o).reshape(n, 32); o += 32 * n
    L = int(np.frombuffer(b, np.int32, 1, o)[0]); o += 4
    sf = np.frombuffer(b, np.float32, L, o); o += 4 * L
    p = O.params(nf)
    okps, odesc, pyr = O.extract(p, img, want_pyr=True)
    assert n == len(okps) and np.array_equal(kps, okps) and np.array_equal(desc, odesc)
    assert np.array_equal(sf.view(np.uint32), np.array(p.scale[:L], np.float32).view(np.uint32))
    po = 0
    for l, (lw, lh) in enumerate(O.level_sizes(p, w, h)):
        cw, ch = np.frombuffer(b, np.int32, 2, o); o += 8
        assert (cw, ch) == (lw, lh)
        got = np.frombuffer(b, np.uint8, (lw + 38) * (lh + 38), o).reshape(lh + 38, lw + 38); o += (lw + 38) * (lh + 38)
        want = O.border(pyr[po:po + lw * lh].reshape(lh, lw), 19); po += lw * lh
        assert np.array_equal(got, want), l
    assert o == len(b)


def _dropin_driver(inp, outp):
    r = subprocess.run([os.path.join(ROOT, "host", "orbx_cpp_matcher_driver"), inp, outp], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_cpp_dropin_matcher(tmp_path):
    """ORB_SLAM2::ORBmatcher (include/ORBmatcher.h, host/ORBmatcher.cc) over stand-in Frame/KeyFrame/MapPoint objects: all
    twelve methods on one scene file against the oracle, and byte for byte against the UNMODIFIED reference ORBmatcher
    (oracle/_ref) run on the same file."""
    import matcher_scene, ref_lib
    subprocess.run(["make", "-C", os.path.join(ROOT, "host"), "-s"], check=True)
    import hashlib, json
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_outputs.json")))["matcher_scenes"]
    got = matcher_scene.check_scene(_dropin_driver, tmp_path)
    assert hashlib.sha256(got).hexdigest() == golden["scene11"]["sha256"]          # the reference's output file, committed as a hash
    for k, name in enumerate(("general21", "general22")):
        g = golden[name]
        d = tmp_path / f"golden{k}"
        d.mkdir()
        out = matcher_scene.check_scene(_dropin_driver, d, g["seed"], g["n1"], g["n2"], general_poses=True)
        assert len(out) == g["bytes"] and hashlib.sha256(out).hexdigest() == g["sha256"], name
    if ref_lib.available():
        rin, rout = str(tmp_path / "m_in.bin"), str(tmp_path / "ref_out.bin")
        assert ref_lib.lib().ref_run_matcher_scene(rin.encode(), rout.encode()) == 0
        assert open(rout, "rb").read() == got
        # a second scene whose projection sections use poses with real rotations, scales and translations (Sim3 1.7 / 0.6,
        # 1-2 degree rotations): no hand-made expectation, the drop-in's output file must equal the reference's
        for k, (seed, n1, n2) in enumerate(((21, 1500, 1300), (22, 900, 1100))):
            d = tmp_path / f"general{k}"
            d.mkdir()
            got = matcher_scene.check_scene(_dropin_driver, d, seed, n1, n2, general_poses=True)
            assert ref_lib.lib().ref_run_matcher_scene(str(d / "m_in.bin").encode(), str(d / "ref_out.bin").encode()) == 0
            want = (d / "ref_out.bin").read_bytes()
            if got != want:
                a, b = matcher_scene._read_all(d / "m_out.bin"), matcher_scene._read_all(d / "ref_out.bin")
                diff = [(i, int((np.frombuffer(x, np.int32) != np.frombuffer(y, np.int32)).sum()) if len(x) == len(y) else -1)
                        for i, (x, y) in enumerate(zip(a, b)) if x != y]
                raise AssertionError(f"scene {seed}: records differing from the reference (index, entries): {diff}")


def pkg_mod():
    from gpu_util import pkg
    return pkg()


def _kfdb_script_bytes(ops):
    """tests/kfdb_scene.py operations in the record format of host/orbx_cpp_kfdb_driver.cc"""
    import struct
    import numpy as np
    code = {"new": 0, "add": 1, "erase": 2, "covis": 3, "loop": 4, "reloc": 5, "clear": 6}
    out = bytearray()
    for op in ops:
        k = op[0]
        if k in ("new", "reloc"):
            w, v = np.asarray(op[2], np.int32), np.asarray(op[3], np.float64)
            out += struct.pack("<iii", code[k], op[1], len(w)) + w.tobytes() + v.tobytes()
        elif k == "covis":
            a = np.asarray(op[2], np.int32)
            out += struct.pack("<iii", 3, op[1], len(a)) + a.tobytes()
        elif k == "loop":
            a = np.asarray(op[2], np.int32)
            out += struct.pack("<iii", 4, op[1], len(a)) + a.tobytes() + struct.pack("<f", op[3])
        else:
            out += struct.pack("<iii", code[k], op[1] if len(op) > 1 else 0, 0)
    return bytes(out)


def test_cpp_dropin_keyframe_database(tmp_path):
    """include/KeyFrameDatabase.h on KeyFrame / Frame stand-ins: add / erase / DetectLoopCandidates /
    DetectRelocalizationCandidates return the keyframes the oracle (and, where oracle/_ref is here, the reference class)
    returns, in the same order, on a generated script and on the hand-made quirk cases."""
    import kfdb_scene as K
    import ref_lib as R
    from test_oracle_kfdb import quirk_script
    subprocess.run(["make", "-C", os.path.join(ROOT, "host"), "-s"], check=True)
    for name, (ops, nw) in {"script": K.make_script(21, nkf=400, spread=120.0, places=5), "quirks": quirk_script()}.items():
        f = tmp_path / (name + ".bin")
        f.write_bytes(_kfdb_script_bytes(ops))
        r = subprocess.run([os.path.join(ROOT, "host", "orbx_cpp_kfdb_driver"), str(f), str(nw)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]
        got = [(ln.split()[0], int(ln.split()[1].rstrip(":")), [int(x) for x in ln.split()[2:]]) for ln in r.stdout.splitlines()]
        o = K.OracleBackend(nw)
        want = K.replay(ops, o)
        o.close()
        assert got == want and len(got) > 5, [(a, b) for a, b in zip(got, want) if a != b][:3]
        if os.path.exists(R.SO):
            rb = K.RefBackend(nw)
            assert K.replay(ops, rb) == got
            rb.close()
