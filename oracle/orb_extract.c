/* TEST INFRASTRUCTURE ONLY (oracle/) -- see orb_oracle.h.
 * CPU restatement of ORBextractor (src/ORBextractor.cc) with the OpenCV primitives
 * replaced by their integer / no-FMA float arithmetic.  Compile with -ffp-contract=off. */
#include "orb_oracle.h"
#include "ref_sincosf.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

static const int kPattern[1024] = {
#include "brief_pattern.inc"
};

/* cvRound(float/double) = round half to even (SSE cvtss2si); default FE mode assumed. */
static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline int cv_round_d(double v) { return (int)lrint(v); }

float orc_sinf(float x) { return ref_sinf(x); }
float orc_cosf(float x) { return ref_cosf(x); }

/* ---------------- ctor tables: src/ORBextractor.cc:471-579 ---------------- */
int orc_params_init(orc_params* p, int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th)
{
    if (nlevels < 1 || nlevels > ORC_MAX_LEVELS || nfeatures < 0) return -1;
    memset(p, 0, sizeof(*p));
    p->nfeatures = nfeatures; p->nlevels = nlevels; p->ini_th = ini_th; p->min_th = min_th;
    p->scale_factor = scale_factor; p->blur_ksize = 9;
    p->scale[0] = 1.0f; p->sigma2[0] = 1.0f;                       /* :488-489 */
    for (int i = 1; i < nlevels; ++i) {                            /* :491-497 */
        p->scale[i] = p->scale[i - 1] * scale_factor;
        p->sigma2[i] = p->scale[i] * p->scale[i];
    }
    for (int i = 0; i < nlevels; ++i) {                            /* :501-505 */
        p->inv_scale[i] = 1.0f / p->scale[i];
        p->inv_sigma2[i] = 1.0f / p->sigma2[i];
    }
    float factor = 1.0f / scale_factor;                            /* :514 */
    float nd = (float)nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels)); /* :516 */
    int sum = 0;
    for (int l = 0; l < nlevels - 1; ++l) {                        /* :521-530 */
        p->nfeat_level[l] = cv_round_f(nd);
        sum += p->nfeat_level[l];
        nd *= factor;
    }
    p->nfeat_level[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0; /* :531 */
    /* umax, :550-578 */
    int v, v0, vmax = (int)floor(ORC_HALF_PATCH * sqrt(2.f) / 2 + 1);
    int vmin = (int)ceil(ORC_HALF_PATCH * sqrt(2.f) / 2);
    const double hp2 = ORC_HALF_PATCH * ORC_HALF_PATCH;
    for (v = 0; v <= vmax; ++v) p->umax[v] = cv_round_d(sqrt(hp2 - v * v));
    for (v = ORC_HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (p->umax[v0] == p->umax[v0 + 1]) ++v0;
        p->umax[v] = v0;
        ++v0;
    }
    return 0;
}

void orc_level_size(const orc_params* p, int level, int w0, int h0, int* w, int* h)
{   /* :1669-1672  Size sz(cvRound((float)cols*scale), cvRound((float)rows*scale)) */
    float s = p->inv_scale[level];
    *w = cv_round_f((float)w0 * s);
    *h = cv_round_f((float)h0 * s);
}

/* ------- cv::resize INTER_LINEAR 8UC1 (called at src/ORBextractor.cc:1686) -------
 * OpenCV imgproc resize.cpp: fixed-point, INTER_RESIZE_COEF_BITS=11. */
static void resize_tab(int dn, int sn, int* ofs, short* c0, short* c1, int clamp_f)
{
    double scale = 1.0 / ((double)dn / (double)sn);
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (clamp_f) { /* x axis: coefficient reset at both ends */
            if (s < 0) { f = 0; s = 0; }
            if (s >= sn - 1) { f = 0; s = sn - 1; }
        }
        ofs[d] = s;
        int a = cv_round_f((1.f - f) * 2048.f), b = cv_round_f(f * 2048.f);
        c0[d] = (short)(a > 32767 ? 32767 : a);
        c1[d] = (short)(b > 32767 ? 32767 : b);
    }
}

void orc_resize_linear(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh, size_t dstride)
{
    int* xo = (int*)malloc(sizeof(int) * (size_t)(dw + dh));
    int* yo = xo + dw;
    short* cx0 = (short*)malloc(sizeof(short) * 2 * (size_t)(dw + dh));
    short *cx1 = cx0 + dw, *cy0 = cx1 + dw, *cy1 = cy0 + dh;
    resize_tab(dw, sw, xo, cx0, cx1, 1);
    resize_tab(dh, sh, yo, cy0, cy1, 0); /* y axis: rows are clipped, beta is kept */
    /* the horizontal pass of a source row is kept while the next destination row still needs it (OpenCV's resize does the
     * same); values are unchanged */
    int* rows = (int*)malloc(sizeof(int) * 2 * (size_t)dw);
    int* xo1 = (int*)malloc(sizeof(int) * (size_t)dw);
    for (int x = 0; x < dw; ++x) xo1[x] = xo[x] + 1 < sw ? xo[x] + 1 : sw - 1;
    int have[2] = {-1, -1};                 /* source row held by rows[0..dw) / rows[dw..2dw) */
    int* T[2] = {rows, rows + dw};
    for (int y = 0; y < dh; ++y) {
        int need[2];
        for (int k = 0; k < 2; ++k) { int sy = yo[y] + k; if (sy < 0) sy = 0; if (sy > sh - 1) sy = sh - 1; need[k] = sy; }
        if (have[1] == need[0] && have[0] != need[0]) { int* t = T[0]; T[0] = T[1]; T[1] = t; have[0] = have[1]; have[1] = -1; }
        for (int k = 0; k < 2; ++k) {
            if (have[k] == need[k]) continue;
            if (k == 1 && have[0] == need[1]) { memcpy(T[1], T[0], sizeof(int) * (size_t)dw); have[1] = need[1]; continue; }
            const uint8_t* S = src + (size_t)need[k] * sstride;
            int* Tk = T[k];
            for (int x = 0; x < dw; ++x) Tk[x] = S[xo[x]] * cx0[x] + S[xo1[x]] * cx1[x];
            have[k] = need[k];
        }
        const int b0 = cy0[y], b1 = cy1[y];
        const int *R0 = T[0], *R1 = T[1];
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < dw; ++x)
            D[x] = (uint8_t)((((b0 * (R0[x] >> 4)) >> 16) + ((b1 * (R1[x] >> 4)) >> 16) + 2) >> 2);
    }
    free(xo1); free(rows); free(cx0); free(xo);
}

/* ------- cv::copyMakeBorder BORDER_REFLECT_101 (:1704,1725) ------- */
static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) { if (i < 0) i = -i; else i = 2 * (n - 1) - i; }
    return i;
}

void orc_border_reflect101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride, int border)
{
    for (int y = -border; y < h + border; ++y) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* D = dst + (size_t)(y + border) * dstride;
        for (int x = -border; x < w + border; ++x) D[x + border] = S[reflect101(x, w)];
    }
}

/* ------- cv::GaussianBlur(Size(k,k), 2, 2, BORDER_REFLECT_101) for CV_8U (:1617-1625) -------
 * OpenCV 4.x fixed-point path: 8.8 kernel (sum 256), (V + 32768) >> 16. */
void orc_gauss_blur(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride, int ksize)
{
    static const int k9[9] = {7, 17, 32, 46, 52, 46, 32, 17, 7};
    static const int k7[7] = {18, 34, 48, 56, 48, 34, 18};
    const int* k = ksize == 9 ? k9 : k7;
    const int r = ksize / 2;
    /* Same arithmetic as before, loops shaped so that the compiler vectorises them (the interior of a row has no border
     * test, the kernel taps are the outer loop of the vertical pass): this file is also the CPU baseline's primitive. */
    uint16_t* H = (uint16_t*)malloc(sizeof(uint16_t) * (size_t)w * (size_t)h);
    uint32_t* A = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)w);
    for (int y = 0; y < h; ++y) {
        const uint8_t* S = src + (size_t)y * sstride;
        uint16_t* Hr = H + (size_t)y * w;
        const int x0 = r < w ? r : w, x1 = w - r > x0 ? w - r : x0;
        for (int x = 0; x < x0; ++x) { int acc = 0; for (int i = 0; i < ksize; ++i) acc += k[i] * S[reflect101(x + i - r, w)]; Hr[x] = (uint16_t)acc; }
        if (ksize == 9)
            for (int x = x0; x < x1; ++x) {
                const uint8_t* q = S + x - 4;
                Hr[x] = (uint16_t)(7 * (q[0] + q[8]) + 17 * (q[1] + q[7]) + 32 * (q[2] + q[6]) + 46 * (q[3] + q[5]) + 52 * q[4]);
            }
        else
            for (int x = x0; x < x1; ++x) { int acc = 0; for (int i = 0; i < ksize; ++i) acc += k[i] * S[x + i - r]; Hr[x] = (uint16_t)acc; }
        for (int x = x1; x < w; ++x) { int acc = 0; for (int i = 0; i < ksize; ++i) acc += k[i] * S[reflect101(x + i - r, w)]; Hr[x] = (uint16_t)acc; }
    }
    for (int y = 0; y < h; ++y) {
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) A[x] = 32768;
        for (int i = 0; i < ksize; ++i) {
            const uint16_t* R = H + (size_t)reflect101(y + i - r, h) * w;
            const uint32_t ki = (uint32_t)k[i];
            for (int x = 0; x < w; ++x) A[x] += ki * R[x];
        }
        for (int x = 0; x < w; ++x) D[x] = (uint8_t)(A[x] >> 16);
    }
    free(A); free(H);
}

/* ------- cv::fastAtan2 scalar (called at :140); no FMA ------- */
float orc_fast_atan2(float y, float x)
{
    static const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    static const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    static const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    static const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/* ------- cv::FAST(roi, kps, th, true): FAST-9/16 (called at :1122,1131) ------- */
static const int kRingX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* OpenCV cornerScore<16>: max(threshold, best arc-of-9 margin) - 1 */
int orc_fast_score(const uint8_t* c, size_t stride, int threshold)
{
    int d[32], v = c[0];
    for (int k = 0; k < 16; ++k) d[k] = v - c[(ptrdiff_t)kRingY[k] * (ptrdiff_t)stride + kRingX[k]];
    for (int k = 16; k < 32; ++k) d[k] = d[k - 16];
    /* min / max over every arc of 9 by doubling (2, 4, 8 consecutive, then the ninth): the same values as the plain double loop */
    int mn2[24], mx2[24], mn4[22], mx4[22];
    for (int k = 0; k < 24; ++k) { mn2[k] = d[k] < d[k + 1] ? d[k] : d[k + 1]; mx2[k] = d[k] > d[k + 1] ? d[k] : d[k + 1]; }
    for (int k = 0; k < 22; ++k) { mn4[k] = mn2[k] < mn2[k + 2] ? mn2[k] : mn2[k + 2]; mx4[k] = mx2[k] > mx2[k + 2] ? mx2[k] : mx2[k + 2]; }
    int best = threshold;
    for (int k = 0; k < 16; ++k) {
        int mn = mn4[k] < mn4[k + 4] ? mn4[k] : mn4[k + 4], mx = mx4[k] > mx4[k + 4] ? mx4[k] : mx4[k + 4];
        if (d[k + 8] < mn) mn = d[k + 8];
        if (d[k + 8] > mx) mx = d[k + 8];
        if (mn > best) best = mn;       /* all 9 darker than centre by >= mn  */
        if (-mx > best) best = -mx;     /* all 9 brighter                     */
    }
    return best - 1;
}

static inline int fast_is_corner(const uint8_t* c, size_t stride, int t)
{
    int v = c[0];
    /* cv::FAST's own early exits (a 9-arc holds ring pixel 0 or 8, and 4 or 12): same result, most pixels leave here */
    {
        const int hi = v + t, lo = v - t;
        const int p0 = c[3 * (ptrdiff_t)stride], p8 = c[-3 * (ptrdiff_t)stride];
        if (!((p0 > hi) | (p0 < lo) | (p8 > hi) | (p8 < lo))) return 0;
        const int p4 = c[3], p12 = c[-3];
        if (!((p4 > hi) | (p4 < lo) | (p12 > hi) | (p12 < lo))) return 0;
    }
    unsigned br = 0, dk = 0; /* ring bitmasks: brighter than v+t / darker than v-t */
    for (int k = 0; k < 16; ++k) {
        int p = c[(ptrdiff_t)kRingY[k] * (ptrdiff_t)stride + kRingX[k]];
        if (p > v + t) br |= 1u << k;
        if (p < v - t) dk |= 1u << k;
    }
    br |= br << 16; dk |= dk << 16;
    unsigned long long B = br | ((unsigned long long)(br & 0xff) << 32), D = dk | ((unsigned long long)(dk & 0xff) << 32);
    unsigned long long rb = B, rd = D;
    for (int i = 1; i < 9; ++i) { rb &= B >> i; rd &= D >> i; }
    return ((rb | rd) & 0xffff) != 0;
}

int orc_fast_roi(const uint8_t* roi, int w, int h, size_t stride, int threshold, int* xys, int cap)
{
    if (w < 7 || h < 7) return 0;
    int* sc = (int*)calloc((size_t)w * (size_t)h, sizeof(int));
    for (int y = 3; y < h - 3; ++y) {
        const uint8_t* S = roi + (size_t)y * stride;
        int x = 3;
#if defined(__SSE2__)
        /* 16 pixels at a time, the way cv::FAST's own SIMD path rejects most of them: a 9-arc holds ring pixel 0 or 8 and ring
         * pixel 4 or 12, so a pixel with both of a pair inside [v - t, v + t] cannot be a corner.  Survivors take the full
         * scalar test below; the result is the same as testing every pixel. */
        const __m128i T = _mm_set1_epi8((char)threshold), Z = _mm_setzero_si128();
        for (; x + 16 <= w - 3; x += 16) {
            const __m128i v = _mm_loadu_si128((const __m128i*)(S + x));
            const __m128i hi = _mm_adds_epu8(v, T), lo = _mm_subs_epu8(v, T);
#define ORC_OUT(p) _mm_or_si128(_mm_subs_epu8((p), hi), _mm_subs_epu8(lo, (p)))   /* non-zero byte: p > v + t or p < v - t */
            const __m128i a = _mm_or_si128(ORC_OUT(_mm_loadu_si128((const __m128i*)(S + x + 3 * (ptrdiff_t)stride))),
                                           ORC_OUT(_mm_loadu_si128((const __m128i*)(S + x - 3 * (ptrdiff_t)stride))));
            const __m128i b = _mm_or_si128(ORC_OUT(_mm_loadu_si128((const __m128i*)(S + x + 3))),
                                           ORC_OUT(_mm_loadu_si128((const __m128i*)(S + x - 3))));
#undef ORC_OUT
            /* byte survives iff a != 0 and b != 0 */
            unsigned m = ~(unsigned)_mm_movemask_epi8(_mm_or_si128(_mm_cmpeq_epi8(a, Z), _mm_cmpeq_epi8(b, Z))) & 0xffffu;
            while (m) {
                const int i = __builtin_ctz(m);
                m &= m - 1;
                const uint8_t* c = S + x + i;
                if (fast_is_corner(c, stride, threshold)) sc[y * w + x + i] = orc_fast_score(c, stride, threshold);
            }
        }
#endif
        for (; x < w - 3; ++x) {
            const uint8_t* c = S + x;
            if (fast_is_corner(c, stride, threshold)) sc[y * w + x] = orc_fast_score(c, stride, threshold);
        }
    }
    int n = 0;
    for (int y = 3; y < h - 3; ++y) {
        const int* row = sc + y * w;
        for (int x = 3; x < w - 3; ++x) {
            const int s = row[x];
            if (s <= 0) continue; /* not a corner (a zero score can never beat its neighbours) */
            const int* r = row + x;
            if (s > r[-1] && s > r[1] && s > r[-w - 1] && s > r[-w] && s > r[-w + 1] && s > r[w - 1] && s > r[w] && s > r[w + 1]) {
                if (n < cap) { xys[3 * n] = x; xys[3 * n + 1] = y; xys[3 * n + 2] = s; }
                ++n;
            }
        }
    }
    free(sc);
    return n;
}

/* ------- per-level cell loop: src/ORBextractor.cc:1058-1153 ------- */
int orc_detect_level(const orc_params* p, const uint8_t* img, int w, int h, size_t stride, orc_keypoint* out, int cap)
{
    const int minBX = ORC_EDGE_THRESHOLD - 3, minBY = minBX;            /* :1061-1062 */
    const int maxBX = w - ORC_EDGE_THRESHOLD + 3, maxBY = h - ORC_EDGE_THRESHOLD + 3; /* :1063-1064 */
    const float W = 30;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);       /* :1076-1077 */
    if (nCols < 1 || nRows < 1) return 0; /* reference divides by zero here (UB); defined as "no keypoints" */
    const int wCell = (int)ceilf(width / nCols), hCell = (int)ceilf(height / nRows); /* :1080-1081 */
    int tmpcap = (wCell + 6) * (hCell + 6);
    int* xys = (int*)malloc(sizeof(int) * 3 * (size_t)tmpcap);
    int n = 0;
    for (int i = 0; i < nRows; ++i) {
        const float iniY = (float)(minBY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;                                  /* :1094 */
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < nCols; ++j) {
            const float iniX = (float)(minBX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;                              /* :1112 */
            if (maxX > maxBX) maxX = (float)maxBX;
            const uint8_t* roi = img + (size_t)(int)iniY * stride + (int)iniX;
            int rw = (int)maxX - (int)iniX, rh = (int)maxY - (int)iniY;
            int m = orc_fast_roi(roi, rw, rh, stride, p->ini_th, xys, tmpcap);       /* :1122 */
            if (m == 0) m = orc_fast_roi(roi, rw, rh, stride, p->min_th, xys, tmpcap); /* :1128-1135 */
            for (int k = 0; k < m; ++k) {                                 /* :1141-1149 */
                if (n < cap) {
                    orc_keypoint* q = &out[n];
                    q->x = (float)xys[3 * k] + (float)(j * wCell);
                    q->y = (float)xys[3 * k + 1] + (float)(i * hCell);
                    q->size = 7.f; q->angle = -1.f; q->response = (float)xys[3 * k + 2];
                    q->octave = 0; q->class_id = -1;
                }
                ++n;
            }
        }
    }
    free(xys);
    return n;
}

/* ------- ExtractorNode / DistributeOctTree: src/ORBextractor.cc:610-682, 696-1042 -------
 * Nodes live in a pool; the std::list is a doubly linked list of pool indices.
 * CANONICAL TIE-BREAK: the reference sorts pair<int,ExtractorNode*> (:935), i.e. equal-size
 * nodes are ordered by heap address (its own comment at :934 flags this as arbitrary).
 * Here the second key is the node's creation sequence number (monotone "address"). */
typedef struct {
    int ulx, uly, brx, bry; /* UL.x, UL.y, BR.x, BR.y (UR/BL are implied) */
    int* keys; int nkeys;
    int no_more;
    int prev, next;         /* list links */
    int seq;
} onode;

typedef struct { onode* n; int cap, cnt, head, tail, size, seq; } olist;

static int ol_new(olist* L)
{
    if (L->cnt == L->cap) { L->cap = L->cap ? 2 * L->cap : 256; L->n = (onode*)realloc(L->n, sizeof(onode) * (size_t)L->cap); }
    onode* q = &L->n[L->cnt];
    memset(q, 0, sizeof(*q)); q->prev = q->next = -1; q->seq = L->seq++;
    return L->cnt++;
}
static void ol_push_front(olist* L, int i)
{
    L->n[i].prev = -1; L->n[i].next = L->head;
    if (L->head >= 0) L->n[L->head].prev = i; else L->tail = i;
    L->head = i; L->size++;
}
static void ol_push_back(olist* L, int i)
{
    L->n[i].next = -1; L->n[i].prev = L->tail;
    if (L->tail >= 0) L->n[L->tail].next = i; else L->head = i;
    L->tail = i; L->size++;
}
static int ol_erase(olist* L, int i) /* returns next */
{
    int p = L->n[i].prev, nx = L->n[i].next;
    if (p >= 0) L->n[p].next = nx; else L->head = nx;
    if (nx >= 0) L->n[nx].prev = p; else L->tail = p;
    L->size--;
    free(L->n[i].keys); L->n[i].keys = 0;
    return nx;
}

/* DivideNode :610-682; children indices returned in c[0..3] = n1..n4 (all allocated) */
static void divide_node(olist* L, int pi, const orc_keypoint* kp, int c[4])
{
    for (int k = 0; k < 4; ++k) c[k] = ol_new(L);
    onode* P = &L->n[pi];
    const int halfX = (int)ceilf((float)(P->brx - P->ulx) / 2);
    const int halfY = (int)ceilf((float)(P->bry - P->uly) / 2);
    const int mx = P->ulx + halfX, my = P->uly + halfY;
    onode *n1 = &L->n[c[0]], *n2 = &L->n[c[1]], *n3 = &L->n[c[2]], *n4 = &L->n[c[3]];
    n1->ulx = P->ulx; n1->uly = P->uly; n1->brx = mx; n1->bry = my;
    n2->ulx = mx; n2->uly = P->uly; n2->brx = P->brx; n2->bry = my;
    n3->ulx = P->ulx; n3->uly = my; n3->brx = mx; n3->bry = P->bry;
    n4->ulx = mx; n4->uly = my; n4->brx = P->brx; n4->bry = P->bry;
    for (int k = 0; k < 4; ++k) L->n[c[k]].keys = (int*)malloc(sizeof(int) * (size_t)(P->nkeys ? P->nkeys : 1));
    for (int i = 0; i < P->nkeys; ++i) {
        const orc_keypoint* q = &kp[P->keys[i]];
        onode* d;
        if (q->x < (float)mx) d = (q->y < (float)my) ? n1 : n3;
        else d = (q->y < (float)my) ? n2 : n4;
        d->keys[d->nkeys++] = P->keys[i];
    }
    for (int k = 0; k < 4; ++k) if (L->n[c[k]].nkeys == 1) L->n[c[k]].no_more = 1;
}

typedef struct { int size, seq, idx; } szptr;
static int szptr_cmp(const void* a, const void* b)
{
    const szptr *x = (const szptr*)a, *y = (const szptr*)b;
    if (x->size != y->size) return x->size < y->size ? -1 : 1;
    return x->seq < y->seq ? -1 : (x->seq > y->seq);
}

int orc_distribute_octree(const orc_keypoint* in, int n, int minX, int maxX, int minY, int maxY, int N, orc_keypoint* out, int cap)
{
    const int nIni = (int)roundf((float)(maxX - minX) / (float)(maxY - minY)); /* :703 */
    if (nIni < 1) return -1; /* reference divides by zero */
    const float hX = (float)(maxX - minX) / (float)nIni;                       /* :706 */
    olist L; memset(&L, 0, sizeof(L)); L.head = L.tail = -1;
    int* ini = (int*)malloc(sizeof(int) * (size_t)nIni);
    for (int i = 0; i < nIni; ++i) {                                           /* :719-738 */
        int q = ol_new(&L);
        L.n[q].ulx = (int)(hX * (float)i); L.n[q].brx = (int)(hX * (float)(i + 1));
        L.n[q].uly = 0; L.n[q].bry = maxY - minY;
        L.n[q].keys = (int*)malloc(sizeof(int) * (size_t)(n ? n : 1));
        ol_push_back(&L, q); ini[i] = q;
    }
    for (int i = 0; i < n; ++i) {                                              /* :742-749 */
        onode* r = &L.n[ini[(int)(in[i].x / hX)]];
        r->keys[r->nkeys++] = i;
    }
    for (int it = L.head; it >= 0;) {                                          /* :756-774 */
        if (L.n[it].nkeys == 1) { L.n[it].no_more = 1; it = L.n[it].next; }
        else if (L.n[it].nkeys == 0) it = ol_erase(&L, it);
        else it = L.n[it].next;
    }
    int finish = 0, nv = 0, vcap = 64;
    szptr* v = (szptr*)malloc(sizeof(szptr) * (size_t)vcap);
#define VPUSH(sz_, idx_) do { if (nv == vcap) { vcap *= 2; v = (szptr*)realloc(v, sizeof(szptr) * (size_t)vcap); } \
        v[nv].size = (sz_); v[nv].seq = L.n[idx_].seq; v[nv].idx = (idx_); ++nv; } while (0)
    while (!finish) {                                                          /* :787 */
        int prevSize = L.size, nToExpand = 0;
        nv = 0;
        for (int it = L.head; it >= 0;) {                                      /* :808-891 */
            if (L.n[it].no_more) { it = L.n[it].next; continue; }
            int c[4];
            divide_node(&L, it, in, c);
            for (int k = 0; k < 4; ++k) {
                if (L.n[c[k]].nkeys > 0) {
                    ol_push_front(&L, c[k]);
                    if (L.n[c[k]].nkeys > 1) { nToExpand++; VPUSH(L.n[c[k]].nkeys, c[k]); }
                } else { free(L.n[c[k]].keys); L.n[c[k]].keys = 0; }
            }
            it = ol_erase(&L, it);
        }
        if (L.size >= N || L.size == prevSize) finish = 1;                     /* :897-903 */
        else if (L.size + nToExpand * 3 > N) {                                 /* :917 */
            while (!finish) {
                prevSize = L.size;
                int np = nv;
                szptr* pv = (szptr*)malloc(sizeof(szptr) * (size_t)(np ? np : 1));
                memcpy(pv, v, sizeof(szptr) * (size_t)np);
                nv = 0;
                qsort(pv, (size_t)np, sizeof(szptr), szptr_cmp);               /* :935 (canonical key) */
                for (int j = np - 1; j >= 0; --j) {                            /* :938 */
                    int c[4];
                    divide_node(&L, pv[j].idx, in, c);
                    for (int k = 0; k < 4; ++k) {
                        if (L.n[c[k]].nkeys > 0) {
                            ol_push_front(&L, c[k]);
                            if (L.n[c[k]].nkeys > 1) VPUSH(L.n[c[k]].nkeys, c[k]);
                        } else { free(L.n[c[k]].keys); L.n[c[k]].keys = 0; }
                    }
                    ol_erase(&L, pv[j].idx);
                    if (L.size >= N) break;                                    /* :991 */
                }
                free(pv);
                if (L.size >= N || L.size == prevSize) finish = 1;             /* :996 */
            }
        }
    }
    int m = 0;
    for (int it = L.head; it >= 0; it = L.n[it].next) {                        /* :1013-1038 */
        const onode* q = &L.n[it];
        int best = q->keys[0];
        float maxR = in[best].response;
        for (int k = 1; k < q->nkeys; ++k)
            if (in[q->keys[k]].response > maxR) { best = q->keys[k]; maxR = in[best].response; }
        if (m < cap) out[m] = in[best];
        ++m;
    }
    for (int i = 0; i < L.cnt; ++i) free(L.n[i].keys);
    free(L.n); free(v); free(ini);
    return m;
}

/* ------- IC_Angle: src/ORBextractor.cc:94-141 ------- */
float orc_ic_angle(const uint8_t* img, size_t stride, int cx, int cy, const int* umax)
{
    int m_01 = 0, m_10 = 0;
    const uint8_t* center = img + (size_t)cy * stride + cx;
    const ptrdiff_t step = (ptrdiff_t)stride;
    for (int u = -ORC_HALF_PATCH; u <= ORC_HALF_PATCH; ++u) m_10 += u * center[u];
    for (int v = 1; v <= ORC_HALF_PATCH; ++v) {
        int v_sum = 0, d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return orc_fast_atan2((float)m_01, (float)m_10);
}

/* ------- computeOrbDescriptor: src/ORBextractor.cc:153-204 ------- */
void orc_brief(const uint8_t* img, size_t stride, int cx, int cy, float angle_deg, uint8_t* desc)
{
    const float factorPI = (float)(3.14159265358979323846 / 180.f); /* :144 */
    float angle = angle_deg * factorPI;
    float a = ref_cosf(angle), b = ref_sinf(angle);
    const uint8_t* center = img + (size_t)cy * stride + cx;
    const ptrdiff_t step = (ptrdiff_t)stride;
    const int* pat = kPattern;
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int k = 0; k < 8; ++k) {
            const int* q = pat + 4 * k;
            float x0 = (float)q[0], y0 = (float)q[1], x1 = (float)q[2], y1 = (float)q[3];
            int t0 = center[cv_round_f(x0 * b + y0 * a) * step + cv_round_f(x0 * a - y0 * b)];
            int t1 = center[cv_round_f(x1 * b + y1 * a) * step + cv_round_f(x1 * a - y1 * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

/* ------- operator(): src/ORBextractor.cc:1543-1658 (+ComputePyramid :1664-1733,
 *         ComputeKeyPointsOctTree :1046-1195) ------- */
int orc_extract(const orc_params* p, const uint8_t* img, int w, int h, size_t stride,
                orc_keypoint* kps, uint8_t* desc, int cap, uint8_t* pyr_out)
{
    if (!img || w <= 0 || h <= 0) return 0;                       /* :1547 empty image -> nothing */
    const int L = p->nlevels;
    int lw[ORC_MAX_LEVELS], lh[ORC_MAX_LEVELS];
    uint8_t* lv[ORC_MAX_LEVELS];
    size_t total = 0;
    for (int l = 0; l < L; ++l) { orc_level_size(p, l, w, h, &lw[l], &lh[l]); if (lw[l] < 1 || lh[l] < 1) return -2; total += (size_t)lw[l] * lh[l]; }
    uint8_t* pyr = (uint8_t*)malloc(total);
    { size_t o = 0; for (int l = 0; l < L; ++l) { lv[l] = pyr + o; o += (size_t)lw[l] * lh[l]; } }
    /* ComputePyramid: level l is resized from level l-1 (chained), :1686 */
    for (int y = 0; y < h; ++y) memcpy(lv[0] + (size_t)y * w, img + (size_t)y * stride, (size_t)w);
    for (int l = 1; l < L; ++l) orc_resize_linear(lv[l - 1], lw[l - 1], lh[l - 1], (size_t)lw[l - 1], lv[l], lw[l], lh[l], (size_t)lw[l]);
    if (pyr_out) memcpy(pyr_out, pyr, total);

    int ntot = 0, status = 0;
    size_t maxpx = (size_t)lw[0] * lh[0];
    orc_keypoint* cand = (orc_keypoint*)malloc(sizeof(orc_keypoint) * maxpx);
    orc_keypoint* sel = (orc_keypoint*)malloc(sizeof(orc_keypoint) * maxpx);
    uint8_t* blur = (uint8_t*)malloc(maxpx);
    for (int l = 0; l < L && status == 0; ++l) {
        const int minB = ORC_EDGE_THRESHOLD - 3;
        const int maxBX = lw[l] - ORC_EDGE_THRESHOLD + 3, maxBY = lh[l] - ORC_EDGE_THRESHOLD + 3;
        int nc = orc_detect_level(p, lv[l], lw[l], lh[l], (size_t)lw[l], cand, (int)maxpx);
        int ns = 0;
        if (nc > 0) {
            ns = orc_distribute_octree(cand, nc, minB, maxBX, minB, maxBY, p->nfeat_level[l], sel, (int)maxpx); /* :1159 */
            if (ns < 0) { status = -3; break; }
        }
        if (ns == 0) continue;                                     /* :1611 */
        const int scaledPatch = (int)((float)ORC_PATCH * p->scale[l]); /* :1168 */
        orc_gauss_blur(lv[l], lw[l], lh[l], (size_t)lw[l], blur, (size_t)lw[l], p->blur_ksize); /* :1617-1625 */
        for (int i = 0; i < ns; ++i) {
            orc_keypoint k = sel[i];
            k.x += (float)minB; k.y += (float)minB; k.octave = l; k.size = (float)scaledPatch; /* :1171-1186 */
            int cx = cv_round_f(k.x), cy = cv_round_f(k.y);
            k.angle = orc_ic_angle(lv[l], (size_t)lw[l], cx, cy, p->umax);                  /* :588-599 */
            if (ntot < cap) {
                orc_brief(blur, (size_t)lw[l], cx, cy, k.angle, desc + (size_t)ntot * 32);  /* :1520-1533 */
                if (l != 0) { k.x *= p->scale[l]; k.y *= p->scale[l]; }                    /* :1642-1651 */
                kps[ntot] = k;
            }
            ++ntot;
        }
    }
    free(blur); free(sel); free(cand); free(pyr);
    return status ? status : ntot;
}

/* cv::cvtColor(src, dst, RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY), 8-bit (call sites src/Tracking.cc:252-278, 313-326,
 * 377-390; third-party OpenCV imgproc, pinned to cv2 4.13: 15-bit fixed point).  channels = 3 or 4, bgr = 1: B first. */
void orc_cvt_gray(const uint8_t* src, int w, int h, size_t sstride, int channels, int bgr, uint8_t* dst, size_t dstride)
{
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const uint8_t* p = src + (size_t)y * sstride + (size_t)x * channels;
            const int r = bgr ? p[2] : p[0], g = p[1], b = bgr ? p[0] : p[2];
            dst[(size_t)y * dstride + x] = (uint8_t)((r * 9798 + g * 19235 + b * 3735 + 16384) >> 15);
        }
}

/* Frame::UndistortKeyPoints, src/Frame.cc:725-781: cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) (third-party
 * OpenCV calib3d, pinned to cv2 4.13: cvUndistortPointsInternal with the default 5 iterations, everything in double, the
 * result stored as float).  K = (fx, fy, cx, cy) and dist (k1, k2, p1, p2[, k3]) are the float entries of mK / mDistCoef. */
void orc_undistort_keypoints(const orc_keypoint* in, int n, float fx_, float fy_, float cx_, float cy_, const float* dist, int ndist,
                             orc_keypoint* out)
{
    if (dist[0] == 0.0f) { for (int i = 0; i < n; ++i) out[i] = in[i]; return; }           /* :727-731 */
    double k[12] = {0};
    for (int i = 0; i < ndist && i < 5; ++i) k[i] = dist[i];
    const double fx = fx_, fy = fy_, cx = cx_, cy = cy_, ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; ++i) {
        out[i] = in[i];
        const double u = in[i].x, v = in[i].y;
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; ++j) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = fx * x + 0.0 * y + cx, yy = 0.0 * x + fy * y + cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        out[i].x = (float)(xx * ww);
        out[i].y = (float)(yy * ww);
    }
}
