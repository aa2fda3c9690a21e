"""CPU: the float matrix algebra of oracle/cvshim (under which the UNMODIFIED reference runs as oracle/_ref) against the
executable OpenCV of this image, cv2 4.13: which products accumulate in float and which in double decides single ulps of the
camera centres and projections that the matchers gate on (src/ORBmatcher.cc:446, :1335, :1760, :1932, src/KeyFrame.cc:102).
host/ORBmatcher.cc (the drop-in) follows the same rules."""
import ctypes
import numpy as np
import pytest
import ref_lib as R

cv2 = pytest.importorskip("cv2")
pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built and /root/reference absent")
f32 = np.float32


def shim(op, A, B=None, C=None):
    A = np.ascontiguousarray(A, f32)
    B = None if B is None else np.ascontiguousarray(B, f32)
    C = None if C is None else np.ascontiguousarray(C, f32)
    out = np.empty(16, f32)
    p = lambda a: None if a is None else a.ctypes.data
    n = R.lib().ref_shim_algebra(op, p(A), A.shape[0], A.shape[1], p(B), 0 if B is None else B.shape[0], 0 if B is None else B.shape[1],
                                 p(C), out.ctypes.data)
    assert n > 0
    return out[:n]


def same(a, b):
    return np.array_equal(np.ascontiguousarray(a, f32).ravel().view(np.uint32), np.ascontiguousarray(b, f32).ravel().view(np.uint32))


def test_products_match_cv2_gemm_bit_for_bit():
    rng = np.random.default_rng(0)
    float_sum_differs = 0
    for _ in range(3000):
        A = rng.normal(size=(3, 3)).astype(f32); B = rng.normal(size=(3, 3)).astype(f32)
        x = (rng.normal(size=(3, 1)) * 5).astype(f32); t = rng.normal(size=(3, 1)).astype(f32)
        assert same(shim(0, A, x), cv2.gemm(A, x, 1, None, 0))                           # Rcw * x3Dw
        assert same(shim(0, A, B), cv2.gemm(A, B, 1, None, 0))                           # sR12 * R2w
        assert same(shim(1, A, x, t), cv2.gemm(A, x, 1, t, 1))                           # Rcw * x3Dw + tcw (one gemm in OpenCV)
        g = cv2.gemm(A, t, 1, None, 0, flags=cv2.GEMM_1_T)
        assert same(shim(2, A, t), g)                                                    # Rcw.t() * tcw: double accumulation
        assert same(shim(3, A, t), cv2.gemm(A, t, -1, None, 0, flags=cv2.GEMM_1_T))      # -Rcw.t() * tcw
        assert same(shim(5, A, t), cv2.gemm(A, t, -1, None, 0))                          # -Rwc * tcw (KeyFrame::SetPose)
        fl = np.array([f32(f32(f32(A[0, r] * t[0, 0]) + f32(A[1, r] * t[1, 0])) + f32(A[2, r] * t[2, 0])) for r in range(3)], f32)
        float_sum_differs += not same(fl, g)
    assert float_sum_differs > 1500          # the float restatement is NOT what OpenCV computes for a transposed operand


def test_norm_dot_and_scalar_division():
    rng = np.random.default_rng(1)
    for _ in range(3000):
        a = rng.normal(size=(3, 1)).astype(f32); b = rng.normal(size=(3, 1)).astype(f32)
        assert R.lib().ref_shim_norm(a.ctypes.data, 3) == cv2.norm(a)                                       # cv::norm(PO)
        assert f32(R.lib().ref_shim_dot(a.ctypes.data, b.ctypes.data, 3)) == cv2.gemm(a, b, 1, None, 0, flags=cv2.GEMM_1_T)[0, 0]
        s = f32(abs(rng.normal()) + 0.3)
        A = rng.normal(size=(3, 3)).astype(f32)
        want = (A * f32(1.0 / np.float64(s))).astype(f32)             # Mat / s = convertTo(alpha = 1/s): element * (float)(1.0 / s)
        assert same(shim(4, A, None, np.array([s], f32)), want)
        assert same(want, cv2.multiply(A, np.full((3, 3), 1.0 / np.float64(s), f32)))   # the same float product through cv2
