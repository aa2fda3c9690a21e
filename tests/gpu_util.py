import importlib
import numpy as np


def pkg():
    return importlib.import_module("orb-slam2-noted_b200")


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def assert_kps_equal(kps, okps, desc, odesc, tag=""):
    assert len(kps) == len(okps), (tag, len(kps), len(okps))
    for f in ("x", "y", "size", "angle", "response"):
        a, b = kps[f].view(np.uint32), okps[f].view(np.uint32)
        assert np.array_equal(a, b), (tag, f, int((a != b).sum()), np.flatnonzero(a != b)[:5])
    assert np.array_equal(kps["octave"], okps["octave"]), tag
    assert np.array_equal(kps["class_id"], okps["class_id"]), tag
    assert np.array_equal(desc, odesc), (tag, int((desc != odesc).any(1).sum()))
