"""CPU: the C-ABI library builds, loads, and exports every symbol include/orbx.h declares."""
import ctypes, os, re
import pytest
from gpu_util import pkg, has_gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for hdr in os.listdir(os.path.join(ROOT, "include")):
        if hdr.endswith(".h") and hdr.startswith("orbx"):
            txt = open(os.path.join(ROOT, "include", hdr)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names |= set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", txt))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    L = ctypes.CDLL(pkg().lib_path())
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"missing export {s}"


def test_no_cpu_fallback_without_device():
    if has_gpu():
        pytest.skip("GPU present")
    P = pkg()
    with pytest.raises(P.OrbxError):
        P.ORBextractor(1000, 1.2, 8, 20, 7)


def test_pattern_tables_identical():
    a = open(os.path.join(ROOT, "oracle", "brief_pattern.inc")).read()
    b = open(os.path.join(ROOT, "orb-slam2-noted_b200", "csrc", "brief_pattern.inc")).read()
    assert a == b


def test_product_never_references_oracle():
    """The product must not import, link, include or call anything under oracle/."""
    bad = ("oracle/", "oracle_lib", "liborb_oracle", "orc_", "orb_oracle.h", "ref_sincosf")
    for dp, _, fs in os.walk(os.path.join(ROOT, "orb-slam2-noted_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".h", ".cc", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                for b in bad:
                    assert b not in txt, (f, b)
