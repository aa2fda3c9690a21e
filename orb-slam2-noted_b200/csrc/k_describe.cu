// K4a + K5 -- IC_Angle (src/ORBextractor.cc:94-141, cv::fastAtan2) and computeOrbDescriptor (:153-204),
// fused: one warp per selected keypoint.  Also performs the final assembly of operator() (:1595-1656):
// level-major concatenation, pt *= scale for level > 0.
//
// Bit-exactness notes (all float ops are explicit round-to-nearest intrinsics, never contracted to FMA,
// because the reference is built for baseline x86-64 without FMA, CMakeLists.txt:16-22):
//  * moments are int32 sums (exact, any reduction order);
//  * fastAtan2 is the scalar OpenCV polynomial evaluated with __fmul_rn/__fadd_rn/__fdiv_rn;
//  * cos/sin follow glibc 2.39's sinf/cosf (__sinf_fma/__cosf_fma variants): double-precision
//    reduction + polynomial with exactly the FMA contraction pattern of that build;
//  * sample coordinates use cvRound = round-half-even = __float2int_rn.
#include "orbx_internal.h"
#include <type_traits>
#include <algorithm>

namespace orbx {

__device__ __align__(16) const float kBriefPatternF[1024] = {   // (x0, y0, x1, y1) per test, src/ORBextractor.cc:209-467
#include "brief_pattern.inc"
};

__device__ __forceinline__ float fastAtan2Dev(float y, float x)
{
    const float k = (float)(180 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * k, p3 = -0.3258083974640975f * k;   // constant-folded in fp32
    const float p5 = 0.1555786518463281f * k, p7 = -0.04432655554792128f * k;
    const float ax = fabsf(x), ay = fabsf(y);
    const float eps = (float)2.2204460492503131e-16; // (float)DBL_EPSILON
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// glibc 2.39 sincosf polynomial; n even: sine, n odd: cosine; neg: coefficients of table[1]
__device__ __forceinline__ float glibcPoly(double x, double x2, int n, bool neg)
{
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5,
                 C3 = -0x1.6c087e89a359dp-10, C4 = 0x1.99343027bf8c3p-16;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __fma_rn(S3, x2, S2);
        const double x7 = __dmul_rn(x3, x2);
        const double s = __fma_rn(x3, S1, x);
        return __double2float_rn(__fma_rn(s1, x7, s));
    }
    const double sg = neg ? -1.0 : 1.0;
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __fma_rn(sg * C4, x2, sg * C3);
    const double c1 = __fma_rn(sg * C1, x2, sg * C0);
    const double x6 = __dmul_rn(x4, x2);
    const double c = __fma_rn(x4, sg * C2, c1);
    return __double2float_rn(__fma_rn(c2, x6, c));
}

__device__ __forceinline__ void glibcSinCosf(float y, float& s, float& c)
{
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
    double x = (double)y;
    if (top < 0x3f4) {
        const double x2 = __dmul_rn(x, x);
        if (top <= 0x397) { s = y; c = 1.0f; return; }
        s = glibcPoly(x, x2, 0, false);
        c = glibcPoly(x, x2, 1, false);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __fma_rn(-(double)n, 0x1.921FB54442D18p0, x);
    const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const double xs = __dmul_rn(x, sign), x2 = __dmul_rn(x, x);
    const bool neg = (n & 2) != 0;
    s = glibcPoly(xs, x2, n, neg);
    c = glibcPoly(xs, x2, n ^ 1, neg);
}

// exported for the exhaustive device-vs-libm test
__global__ void sincosfProbeKernel(uint32_t firstBits, uint32_t count, float* s, float* c)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    glibcSinCosf(__uint_as_float(firstBits + i), s[i], c[i]);
}

__global__ void atan2ProbeKernel(const float* yx, int n, float* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = fastAtan2Dev(yx[2 * i], yx[2 * i + 1]);
}

constexpr int kDescWarps = 8;
constexpr int kDepth = 2;                                                       // patches in flight per warp (3 measured no faster)
constexpr int kKpPerWarpMax = 32;                                               // consecutive keypoints per warp (lane k owns keypoint k), patches double-buffered
constexpr int kPatchR = 18, kPatchRows = 2 * kPatchR + 1, kPatchPitchW = 16;   // blurred patch: 37 rows x 64 bytes (four 16-byte chunks)
constexpr int kRawRows = 2 * kHalfPatch + 1, kRawPitchW = 12;                  // raw patch: 31 rows x 48 bytes (three 16-byte chunks)
constexpr int kPatchWords = kPatchRows * kPatchPitchW, kRawWords = kRawRows * kRawPitchW;

__device__ __forceinline__ void cpAsync16(uint32_t dst, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}

__global__ void __launch_bounds__(kDescWarps * 32, 3)
describeKernel(const __grid_constant__ GeomSet G, const uint8_t* __restrict__ lvl0, size_t lvl0Pitch, size_t lvl0Frame,
               const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blur,
               const orbx_keypoint* __restrict__ selAll, const int* __restrict__ selCount,
               orbx_keypoint* __restrict__ outKp, uint8_t* __restrict__ outDesc, int cap, int* __restrict__ outN,
               int* __restrict__ err, int kpPerWarp, const int* __restrict__ frameOff)
{
    extern __shared__ __align__(16) uint32_t sBufAll[];                      // [kDescWarps][2][kPatchWords]; pass A uses it as [2][kRawWords]
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int frame = blockIdx.y + G.frame0;

    // level-major output index = running count over the levels (final assembly of operator(), :1595-1656)
    const int cnt = lane < G.nlevels ? selCount[frame * G.nlevels + lane] : 0;
    int incl = cnt;
    #pragma unroll
    for (int o = 1; o < kMaxLevels; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const int total = __shfl_sync(0xffffffffu, incl, G.nlevels - 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) outN[frame] = min(total, cap);
    const int i0 = (blockIdx.x * kDescWarps + wrp) * kpPerWarp;
    // first output row of this frame: its padded block, or (ORBX_OUT_DENSE) right behind the previous frame's records
    const size_t outBase = frameOff ? (size_t)frameOff[frame] : (size_t)frame * cap;
    if (i0 >= total) return;
    if (total > cap) { if (lane == 0) atomicExch(err, 5); if (i0 >= cap) return; }
    const int nkp = min(kpPerWarp, min(total, cap) - i0);

    // lane k < nkp owns keypoint i0 + k: its level, its record
    int myLevel = 0; orbx_keypoint kp = {};
    {
        const int i = i0 + min(lane, nkp - 1);
        #pragma unroll
        for (int l = 0; l < kMaxLevels; ++l) myLevel += (l < G.nlevels - 1 && __shfl_sync(0xffffffffu, incl, l) <= i) ? 1 : 0;
        const int before = __shfl_sync(0xffffffffu, incl - cnt, myLevel);
        if (lane < nkp) kp = selAll[(size_t)frame * G.selPerFrame + G.lv[myLevel].outBase + (i - before)];
    }

    // ---- two passes over the warp's keypoints, each streaming its patches through a double buffer of 16-byte async
    //      copies (rows of every plane are 16-byte aligned; a patch is fetched as the chunks that cover it):
    //      pass A  31 x 31 raw patch (3 chunks per row) -> IC_Angle moments, lane k keeps the moments of keypoint k;
    //              then fastAtan2 + sinf/cosf run ONCE, lane-parallel over the warp's keypoints
    //      pass B  37 x 37 blurred patch (rotated pattern points stay within +-18; 4 chunks per row) -> steered BRIEF ----
    uint32_t* const bufBase = sBufAll + wrp * (kDepth * kPatchWords);
    auto issueRaw = [&](int k) {
        const int level = __shfl_sync(0xffffffffu, myLevel, k);
        const int cx = __float2int_rn(__shfl_sync(0xffffffffu, kp.x, k)), cy = __float2int_rn(__shfl_sync(0xffffffffu, kp.y, k));
        const LevelGeom& g = G.lv[level];
        uint32_t* buf = bufBase + (k % kDepth) * kRawWords;
        const uint8_t* img; int pitch;
        if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = (int)lvl0Pitch; }
        else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = g.pitch; }
        const int ra0 = (cx - kHalfPatch) & ~15;
        if (((((size_t)img) | (size_t)pitch) & 15) == 0) {
            const int lr = lane / 3, lk = lane - lr * 3;            // 10 rows x 3 chunks per step
            const bool ok = lane < 30 && ra0 + 16 * lk < pitch;     // a chunk beyond the padded row is never used
            const uint8_t* src = img + (size_t)(cy - kHalfPatch + lr) * pitch + ra0 + 16 * lk;
            uint32_t dst = (uint32_t)__cvta_generic_to_shared(buf + lr * kRawPitchW + 4 * lk);
            const int step = 10 * pitch;
            #pragma unroll
            for (int t = 0; t < (kRawRows + 9) / 10; ++t) {
                if (ok && 10 * t + lr < kRawRows) cpAsync16(dst, src);
                src += step; dst += 10 * kRawPitchW * 4;
            }
        } else {                                                     // caller image without 16-byte rows (level 0 only): byte loads
            uint8_t* rb = reinterpret_cast<uint8_t*>(buf);
            const int offr = (cx - kHalfPatch) - ra0;
            for (int i = lane; i < kRawRows * 31; i += 32) {
                const int r = i / 31, c = i - r * 31;
                rb[r * (kRawPitchW * 4) + offr + c] = img[(size_t)(cy - kHalfPatch + r) * pitch + cx - kHalfPatch + c];
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto issueBlur = [&](int k) {
        const int level = __shfl_sync(0xffffffffu, myLevel, k);
        const int cx = __float2int_rn(__shfl_sync(0xffffffffu, kp.x, k)), cy = __float2int_rn(__shfl_sync(0xffffffffu, kp.y, k));
        const LevelGeom& g = G.lv[level];
        uint32_t* buf = bufBase + (k % kDepth) * kPatchWords;
        const int bp = g.pitch;
        const int a0 = (cx - kPatchR) & ~15;                        // aligned first column of the blurred patch
        const int lr = lane >> 2, lk = lane & 3;                    // 8 rows x 4 chunks per step
        const bool ok = a0 + 16 * lk < bp;
        const uint8_t* src = blur + g.blurOff + ((size_t)frame * g.h + (cy - kPatchR + lr)) * bp + a0 + 16 * lk;
        uint32_t dst = (uint32_t)__cvta_generic_to_shared(buf + lr * kPatchPitchW + 4 * lk);
        const int step = 8 * bp;
        #pragma unroll
        for (int t = 0; t < (kPatchRows + 7) / 8; ++t) {
            if (ok && 8 * t + lr < kPatchRows) cpAsync16(dst, src);
            src += step; dst += 8 * kPatchPitchW * 4;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    // ---- pass A: IC_Angle.  lane = row v of the disc: the row's 31 bytes are aligned into 8 words, bytes outside
    //      |u| <= umax[|v|] are masked off, and IDP.4A gives the row sum (m01 = v * sum) and the sum of u * p. ----
    int myM10 = 0, myM01 = 0;
    {
        uint32_t icMask[8];
        const int v = lane - kHalfPatch, d = lane < 2 * kHalfPatch + 1 ? G.umax[v < 0 ? -v : v] : -1;
        const int lo = kHalfPatch - d, hi = kHalfPatch + d;            // inclusive byte range (empty when d < 0)
        #pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint32_t m = 0;
            #pragma unroll
            for (int b2 = 0; b2 < 4; ++b2) if (4 * j + b2 >= lo && 4 * j + b2 <= hi) m |= 0xffu << (8 * b2);
            icMask[j] = m;
        }
        #pragma unroll
        for (int d = 0; d < kDepth; ++d) { if (d < nkp) issueRaw(d); else asm volatile("cp.async.commit_group;" ::: "memory"); }
        for (int k = 0; k < nkp; ++k) {
            asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");   // one group per keypoint, kDepth in flight
            __syncwarp();
            const int cx = __float2int_rn(__shfl_sync(0xffffffffu, kp.x, k));
            const uint32_t* buf = bufBase + (k % kDepth) * kRawWords;
            int m10 = 0, m01 = 0;
            if (lane < 2 * kHalfPatch + 1) {
                // the row's 12 words as three 128-bit loads: with a pitch of 12 words the 8 lanes of a quarter warp cover all 32
                // banks (32-bit loads of one word column would hit only 8 banks: 4-way conflicts)
                const uint4* row4 = reinterpret_cast<const uint4*>(buf + lane * kRawPitchW);
                const uint4 q0 = row4[0], q1 = row4[1], q2 = row4[2];
                const uint32_t w[12] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w};
                const uint32_t sh = 8u * (uint32_t)((cx - kHalfPatch) & 3);                         // byte of u = -15 inside its word
                int cs = 0;
                auto rowSums = [&](auto oc) {                                                       // oc = word holding u = -15
                    constexpr int o = decltype(oc)::value;
                    #pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint32_t x = __funnelshift_r(w[o + j], w[o + j + 1], sh) & icMask[j];
                        // signed weights u = 4j-15 .. 4j-12 of the word's four bytes
                        const int uw = (int)(((uint32_t)(uint8_t)(4 * j - 15)) | ((uint32_t)(uint8_t)(4 * j - 14) << 8) |
                                             ((uint32_t)(uint8_t)(4 * j - 13) << 16) | ((uint32_t)(uint8_t)(4 * j - 12) << 24));
                        cs = (int)__dp4a(x, 0x01010101u, (uint32_t)cs);
                        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(m10) : "r"(x), "r"(uw));
                    }
                };
                switch (((cx - kHalfPatch) & 15) >> 2) {                                            // warp-uniform
                    case 0: rowSums(std::integral_constant<int, 0>()); break;
                    case 1: rowSums(std::integral_constant<int, 1>()); break;
                    case 2: rowSums(std::integral_constant<int, 2>()); break;
                    default: rowSums(std::integral_constant<int, 3>()); break;
                }
                m01 = (lane - kHalfPatch) * cs;
            }
            m10 = __reduce_add_sync(0xffffffffu, m10);      // REDUX: exact integer sums, any order
            m01 = __reduce_add_sync(0xffffffffu, m01);
            if (lane == k) { myM10 = m10; myM01 = m01; }
            __syncwarp();                                    // everyone is done with this buffer
            if (k + kDepth < nkp) issueRaw(k + kDepth); else asm volatile("cp.async.commit_group;" ::: "memory");
        }
    }
    // orientation and its sine / cosine, one keypoint per lane (:140, :158)
    const float myAngle = fastAtan2Dev((float)myM01, (float)myM10);
    float myCos, mySin;
    {
        const float factorPI = (float)(3.14159265358979323846 / 180.f);
        glibcSinCosf(__fmul_rn(myAngle, factorPI), mySin, myCos);
    }

    // ---- pass B: steered BRIEF on the blurred patch: lane = descriptor byte; the lane's 8 test pairs of the pattern
    //      stay in registers for all keypoints of the warp ----
    float4 pat[8];
    #pragma unroll
    for (int j = 0; j < 8; ++j) pat[j] = __ldg(reinterpret_cast<const float4*>(kBriefPatternF) + lane * 8 + j);
    #pragma unroll
    for (int d = 0; d < kDepth; ++d) { if (d < nkp) issueBlur(d); else asm volatile("cp.async.commit_group;" ::: "memory"); }
    for (int k = 0; k < nkp; ++k) {
        asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");
        __syncwarp();
        const int cx = __float2int_rn(__shfl_sync(0xffffffffu, kp.x, k));
        const float a = __shfl_sync(0xffffffffu, myCos, k), b = __shfl_sync(0xffffffffu, mySin, k);
        const uint32_t* buf = bufBase + (k % kDepth) * kPatchWords;
        const int offc = (cx - kPatchR) - ((cx - kPatchR) & ~15) + kPatchR;                // byte column of the patch centre
        const uint8_t* pc = reinterpret_cast<const uint8_t*>(buf) + kPatchR * (kPatchPitchW * 4) + offc;
        // cvRound without the conversion unit: x + 1.5 * 2^23 rounds x (|x| < 2^22) to the nearest integer, ties to even,
        // and leaves it in the low mantissa bits; the constant part of (row * pitch + col) is folded into the base address
        constexpr float kMagic = 12582912.0f;
        constexpr uint32_t kMagicBits = 0x4B400000u, kPitchB = kPatchPitchW * 4;
        const uint32_t pcm = (uint32_t)__cvta_generic_to_shared(pc) - kMagicBits * (kPitchB + 1u);
        auto sample = [&](float fx, float fy) -> uint32_t {          // pattern point (fx, fy) rotated by the keypoint angle
            const float r = __fadd_rn(__fadd_rn(__fmul_rn(fx, b), __fmul_rn(fy, a)), kMagic);
            const float c = __fadd_rn(__fsub_rn(__fmul_rn(fx, a), __fmul_rn(fy, b)), kMagic);
            uint32_t v;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(pcm + __float_as_uint(r) * kPitchB + __float_as_uint(c)));
            return v;
        };
        int val = 0;
        #pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 q = pat[j];
            const uint32_t t0 = sample(q.x, q.y), t1 = sample(q.z, q.w);
            val |= (t0 < t1) << j;
        }
        outDesc[(outBase + (i0 + k)) * 32 + lane] = (uint8_t)val;
        __syncwarp();                                              // everyone is done with this buffer
        if (k + kDepth < nkp) issueBlur(k + kDepth); else asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (lane < nkp) {
        kp.angle = myAngle;
        if (myLevel != 0) { const float sc = G.lv[myLevel].scale; kp.x = __fmul_rn(kp.x, sc); kp.y = __fmul_rn(kp.y, sc); }
        outKp[outBase + i0 + lane] = kp;
    }
}

// Frame::UndistortKeyPoints (src/Frame.cc:725-781) = cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on the
// keypoint coordinates (third-party OpenCV calib3d; pinned to cv2 4.13): normalise with K, 5 fixed-point iterations of the
// radial / tangential model in double precision, re-project with K, round to float.  Every double operation is an explicit
// round-to-nearest intrinsic in OpenCV's order (the x86-64 baseline build has no FMA).
struct UndistortP { double fx, fy, cx, cy, k[5]; };
__global__ void undistortKernel(const orbx_keypoint* __restrict__ in, int n, const UndistortP P, orbx_keypoint* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orbx_keypoint kp = in[i];
    const double ifx = __ddiv_rn(1.0, P.fx), ify = __ddiv_rn(1.0, P.fy);
    double x = __dmul_rn(__dsub_rn((double)kp.x, P.cx), ifx), y = __dmul_rn(__dsub_rn((double)kp.y, P.cy), ify);
    const double x0 = x, y0 = y;
    const double k1 = P.k[0], k2 = P.k[1], p1 = P.k[2], p2 = P.k[3], k3 = P.k[4];
    bool bad = false;
    for (int j = 0; j < 5 && !bad; ++j) {
        const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
        // (1 + ((k[7]*r2 + k[6])*r2 + k[5])*r2) / (1 + ((k[4]*r2 + k[1])*r2 + k[0])*r2) with k[5..7] = 0
        const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(0.0, r2), 0.0), r2), 0.0), r2));
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k3, r2), k2), r2), k1), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) { x = x0; y = y0; bad = true; break; }
        // deltaX = 2*p1*x*y + p2*(r2 + 2*x*x) (+ 0*r2 + 0*r2*r2), deltaY = p1*(r2 + 2*y*y) + 2*p2*x*y (+ ...)
        const double dX = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, p1), x), y),
                                                        __dmul_rn(p2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x)))),
                                              __dmul_rn(0.0, r2)), __dmul_rn(__dmul_rn(0.0, r2), r2));
        const double dY = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(p1, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                                        __dmul_rn(__dmul_rn(__dmul_rn(2.0, p2), x), y)),
                                              __dmul_rn(0.0, r2)), __dmul_rn(__dmul_rn(0.0, r2), r2));
        x = __dmul_rn(__dsub_rn(x0, dX), icdist);
        y = __dmul_rn(__dsub_rn(y0, dY), icdist);
    }
    // RR = K: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1/(0*x + 0*y + 1)
    const double xx = __dadd_rn(__dadd_rn(__dmul_rn(P.fx, x), __dmul_rn(0.0, y)), P.cx);
    const double yy = __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(P.fy, y)), P.cy);
    const double ww = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(0.0, y)), 1.0));
    kp.x = __double2float_rn(__dmul_rn(xx, ww));
    kp.y = __double2float_rn(__dmul_rn(yy, ww));
    out[i] = kp;
}

// ORBX_OUT_DENSE: off[f] = sum over the frames before f of min(keypoints selected, cap); off[frames] = total.  One CTA.
__global__ void __launch_bounds__(1024) frameOffsetsKernel(const int* __restrict__ selCount, int nlevels, int frames, int cap, int* __restrict__ off)
{
    __shared__ int warpSum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int base = 0; base < frames; base += 1024) {
        const int f = base + threadIdx.x;
        int v = 0;
        if (f < frames) { for (int l = 0; l < nlevels; ++l) v += selCount[f * nlevels + l]; v = min(v, cap); }
        int s = v;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += t; }
        if (lane == 31) warpSum[w] = s;
        __syncthreads();
        if (w == 0) {
            int t = warpSum[lane];
            #pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += u; }
            warpSum[lane] = t;
        }
        __syncthreads();
        const int before = carry + (w ? warpSum[w - 1] : 0) + s - v;
        if (f < frames) off[f] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[frames] = carry;
}

int launchDescribe(Ctx* c, int frame0, int batch, orbx_keypoint* dKps, uint8_t* dDesc, int capPerFrame, int* dN, bool dense)
{
    GeomSet g = c->geom;
    g.frame0 = frame0;
    if (dense) {
        if (frame0 != 0) return ORBX_E_INVALID;                  // dense output is one pass over the whole batch
        if (!c->dFrameOff) ORBX_CUDA(c, cudaMalloc(&c->dFrameOff, sizeof(int) * (size_t)(c->maxBatch + 1)));
    }
    stageBegin(c, ST_DESCRIBE);
    if (dense) frameOffsetsKernel<<<1, 1024, 0, c->stream>>>(c->dSelCount, g.nlevels, batch, capPerFrame, c->dFrameOff);
    // keypoints per warp: as many as keeps every SM busy (the pattern registers and the count prefix are per warp)
    const int slots = std::min(g.selPerFrame, capPerFrame);
    int kpw = kKpPerWarpMax;
    int smCount = 148;
    cudaDeviceGetAttribute(&smCount, cudaDevAttrMultiProcessorCount, c->device);
    while (kpw > 2 && (long long)batch * ((slots + kpw - 1) / kpw) < (long long)smCount * 2 * kDescWarps) kpw >>= 1;
    const int perBlock = kDescWarps * kpw;
    dim3 grid((slots + perBlock - 1) / perBlock, batch), block(kDescWarps * 32);
    const size_t smem = sizeof(uint32_t) * (size_t)kDescWarps * kDepth * kPatchWords;
    if (smem > c->smemCfg[SMEM_DESCRIBE]) {
        ORBX_CUDA(c, cudaFuncSetAttribute(describeKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        c->smemCfg[SMEM_DESCRIBE] = smem;
    }
    describeKernel<<<grid, block, smem, c->stream>>>(g, c->lvl0, c->lvl0Pitch, c->lvl0FrameStride, c->dPyr, c->dBlur,
                                                 c->dSel, c->dSelCount, dKps, dDesc, capPerFrame, dN, c->dErr, kpw,
                                                 dense ? c->dFrameOff : nullptr);
    stageEnd(c, ST_DESCRIBE, dense ? 2 : 1);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx

// dist: mDistCoef (k1, k2, p1, p2[, k3]) as floats, ndist = 4 or 5; dist[0] == 0 copies the keypoints (src/Frame.cc:727-731).
extern "C" int orbx_undistort_keypoints(orbx_ctx* h, const orbx_keypoint* keys, int n, float fx, float fy, float cx, float cy,
                                        const float* dist, int ndist, int flags, orbx_keypoint* out)
{
    if (!h || n < 0 || !dist || ndist < 4 || ndist > 5 || (n > 0 && (!keys || !out))) return ORBX_E_INVALID;
    if (n == 0) return ORBX_OK;
    orbx::Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const size_t bytes = sizeof(orbx_keypoint) * (size_t)n;
    if (dist[0] == 0.0f) {
        const cudaMemcpyKind kind = inDev ? (outDev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost)
                                          : (outDev ? cudaMemcpyHostToDevice : cudaMemcpyHostToHost);
        if (keys != out) ORBX_CUDA(c, cudaMemcpyAsync(out, keys, bytes, kind, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        return ORBX_OK;
    }
    int rc = orbx::ensureScratch(c, 2 * ((bytes + 255) & ~(size_t)255));
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_keypoint* dIn = keys; orbx_keypoint* dOut = out;
    if (!inDev) { ORBX_CUDA(c, cudaMemcpyAsync(S, keys, bytes, cudaMemcpyHostToDevice, c->stream)); dIn = (const orbx_keypoint*)S; }
    if (!outDev) dOut = (orbx_keypoint*)(S + ((bytes + 255) & ~(size_t)255));
    orbx::UndistortP P;
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
    for (int i = 0; i < 5; ++i) P.k[i] = i < ndist ? (double)dist[i] : 0.0;
    orbx::undistortKernel<<<(n + 255) / 256, 256, 0, c->stream>>>(dIn, n, P, dOut);
    c->kernelLaunches += 1;
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) ORBX_CUDA(c, cudaMemcpyAsync(out, dOut, bytes, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}

// ---- test probes (C ABI, declared in include/orbx_test.h) ----
extern "C" int orbx_test_sincosf(uint32_t first_bits, uint32_t count, float* h_sin, float* h_cos)
{
    float *ds = nullptr, *dc = nullptr;
    if (cudaMalloc(&ds, sizeof(float) * (size_t)count) != cudaSuccess) return ORBX_E_CUDA;
    if (cudaMalloc(&dc, sizeof(float) * (size_t)count) != cudaSuccess) { cudaFree(ds); return ORBX_E_CUDA; }
    orbx::sincosfProbeKernel<<<(count + 255) / 256, 256>>>(first_bits, count, ds, dc);
    cudaError_t e = cudaMemcpy(h_sin, ds, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(h_cos, dc, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost);
    cudaFree(ds); cudaFree(dc);
    return e == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbx_test_fast_atan2(const float* h_yx, int n, float* h_out)
{
    float *d = nullptr, *o = nullptr;
    if (cudaMalloc(&d, sizeof(float) * 2 * (size_t)n) != cudaSuccess) return ORBX_E_CUDA;
    if (cudaMalloc(&o, sizeof(float) * (size_t)n) != cudaSuccess) { cudaFree(d); return ORBX_E_CUDA; }
    cudaMemcpy(d, h_yx, sizeof(float) * 2 * (size_t)n, cudaMemcpyHostToDevice);
    orbx::atan2ProbeKernel<<<(n + 255) / 256, 256>>>(d, n, o);
    cudaError_t e = cudaMemcpy(h_out, o, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost);
    cudaFree(d); cudaFree(o);
    return e == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}
