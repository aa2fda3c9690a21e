"""CPU: exhaustive pin of the glibc-2.39 sinf/cosf restatement (oracle/ref_sincosf.h) against the
host libm over every float in [0, 6.5] (1.09e9 values, a few seconds on 8 threads)."""
import json, os, subprocess
import oracle_lib as O


def test_sincosf_exhaustive_vs_host_libm():
    O.build()
    exe = os.path.join(O.ODIR, "check_sincosf")
    r = subprocess.run([exe, "0", "6.5", str(os.cpu_count() or 4)], capture_output=True, text=True)
    rep = json.loads(r.stdout)
    if not rep["glibc"].startswith("2.39"):
        import pytest
        pytest.skip("restatement is pinned to glibc 2.39; host has " + rep["glibc"])
    assert rep["bad_sin"] == 0 and rep["bad_cos"] == 0, rep
    assert rep["n"] > 1_080_000_000
