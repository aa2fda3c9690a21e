/* TEST INFRASTRUCTURE ONLY (oracle/) -- see orb_oracle.h.
 * CPU restatement of the Hamming side of the reference: ORBmatcher (src/ORBmatcher.cc),
 * Frame::ComputeStereoMatches / GetFeaturesInArea / AssignFeaturesToGrid (src/Frame.cc) and the
 * DBoW2 tree descent (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h, FORB.cpp, FeatureVector.cpp).
 * SLAM objects (MapPoint*, KeyFrame*, cv::Mat) are flattened to plain arrays; every loop keeps the
 * reference's iteration order, tie rules and float/double mix.  Compile with -ffp-contract=off. */
#include "orb_oracle.h"
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define TH_HIGH 100     /* src/ORBmatcher.cc:56 */
#define TH_LOW 50       /* :57 */
#define HISTO_LENGTH 30 /* :58 */

/* DescriptorDistance: src/ORBmatcher.cc:2123-2143 (8 x 32-bit bit-hack popcount); FORB::distance
 * (Thirdparty/DBoW2/DBoW2/FORB.cpp:81-101) is the same arithmetic. */
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t x, y;
        memcpy(&x, a + 4 * i, 4); memcpy(&y, b + 4 * i, 4);
        uint32_t v = x ^ y;
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
    }
    return dist;
}

/* ComputeThreeMaxima: src/ORBmatcher.cc:2076-2118 (on bin counts) */
void orc_three_maxima(const int* count, int L, int* ind1, int* ind2, int* ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = count[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if ((float)max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if ((float)max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

/* rotation bin: src/ORBmatcher.cc:364-371 (factor = HISTO_LENGTH/360.0f in this fork, :275) */
static int rot_bin(float a1, float a2)
{
    const float factor = HISTO_LENGTH / 360.0f;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)roundf(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

typedef struct { int* v; int n, cap; } ivec;
static void ipush(ivec* a, int x)
{
    if (a->n == a->cap) { a->cap = a->cap ? 2 * a->cap : 64; a->v = (int*)realloc(a->v, sizeof(int) * (size_t)a->cap); }
    a->v[a->n++] = x;
}

/* ---------------- DBoW2 ---------------- */
/* TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), TemplatedVocabulary.h:1255-1306 */
void orc_bow_transform(const orc_vocab* V, const uint8_t* desc, int n, int levelsup, int32_t* word, float* weight, int32_t* node)
{
    const int nid_level = V->L - levelsup;
    for (int f = 0; f < n; ++f) {
        const uint8_t* d = desc + 32 * (size_t)f;
        int nid = 0, final_id = 0, level = 0;
        do {
            ++level;
            const int c0 = V->child_off[final_id], c1 = V->child_off[final_id + 1];
            final_id = V->child[c0];
            double best = (double)orc_descriptor_distance(d, V->desc + 32 * (size_t)final_id);
            for (int c = c0 + 1; c < c1; ++c) {
                const int id = V->child[c];
                const double dd = (double)orc_descriptor_distance(d, V->desc + 32 * (size_t)id);
                if (dd < best) { best = dd; final_id = id; }
            }
            if (level == nid_level) nid = final_id;
        } while (V->child_off[final_id + 1] > V->child_off[final_id]); /* !isLeaf() */
        word[f] = V->word_id[final_id];
        weight[f] = V->weight[final_id];
        node[f] = nid;
    }
}

/* Best / second-best state of the inner loops at src/ORBmatcher.cc:325-345 (strict <, first train wins ties), for every
 * query of every segment against the trains of the same segment; t_valid models the `continue`s at :323,:822-826.
 * Rows are reported in q_idx list order; an empty candidate set leaves (-1, 256, 256). */
void orc_hamming_best2(const uint8_t* q, const uint8_t* t, const int32_t* q_seg, const int32_t* q_idx, const int32_t* t_seg,
                       const int32_t* t_idx, int nseg, const uint8_t* t_valid, int32_t* best_idx, int32_t* best_dist,
                       int32_t* second_dist)
{
    for (int s = 0; s < nseg; ++s)
        for (int a = q_seg[s]; a < q_seg[s + 1]; ++a) {
            const int iq = q_idx ? q_idx[a] : a;
            int best1 = 256, best2 = 256, bi = -1;
            for (int b = t_seg[s]; b < t_seg[s + 1]; ++b) {
                const int it = t_idx ? t_idx[b] : b;
                if (t_valid && !t_valid[it]) continue;
                const int dist = orc_descriptor_distance(q + 32 * (size_t)iq, t + 32 * (size_t)it);
                if (dist < best1) { best2 = best1; best1 = dist; bi = it; }
                else if (dist < best2) best2 = dist;
            }
            best_idx[a] = bi; best_dist[a] = best1; second_dist[a] = best2;
        }
}

/* FeatureVector::addFeature (FeatureVector.cpp:37-54) for every feature with weight > 0
 * (TemplatedVocabulary.h:1184-1189): std::map order = ascending node id, features ascending.
 * Returns number of nodes; node_id[nn], node_off[nn+1], feat[n_kept]. */
int orc_feature_vector(const int32_t* node, const float* weight, int n, int32_t* node_id, int32_t* node_off, int32_t* feat)
{
    int* order = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    int m = 0;
    for (int i = 0; i < n; ++i) if (weight[i] > 0) order[m++] = i;
    /* stable insertion by node id (n is small in tests; oracle clarity over speed) -> merge sort */
    int* tmp = (int*)malloc(sizeof(int) * (size_t)(m > 0 ? m : 1));
    for (int w = 1; w < m; w *= 2) {
        for (int lo = 0; lo < m; lo += 2 * w) {
            int mid = lo + w < m ? lo + w : m, hi = lo + 2 * w < m ? lo + 2 * w : m;
            int a = lo, b = mid, k = lo;
            while (a < mid && b < hi) tmp[k++] = node[order[b]] < node[order[a]] ? order[b++] : order[a++];
            while (a < mid) tmp[k++] = order[a++];
            while (b < hi) tmp[k++] = order[b++];
        }
        memcpy(order, tmp, sizeof(int) * (size_t)m);
    }
    int nn = 0;
    for (int i = 0; i < m; ++i) {
        if (i == 0 || node[order[i]] != node[order[i - 1]]) { node_id[nn] = node[order[i]]; node_off[nn] = i; ++nn; }
        feat[i] = order[i];
    }
    node_off[nn] = m;
    free(order); free(tmp);
    return nn;
}

/* merge-join of two FeatureVectors (the while/lower_bound loops, e.g. src/ORBmatcher.cc:285-390) */
#define FOR_SHARED_NODES(A, B, ia, ib) \
    for (int ia = 0, ib = 0; ia < (A)->nnodes && ib < (B)->nnodes;) \
        if ((A)->node_id[ia] < (B)->node_id[ib]) ++ia; else if ((A)->node_id[ia] > (B)->node_id[ib]) ++ib; else

/* SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches): src/ORBmatcher.cc:255-419.
 * kf->valid[i] = KF feature i has a good MapPoint.  match_f[F.N] receives the KF feature index whose
 * MapPoint was assigned to F feature j (vpMapPointMatches[j]), or -1. */
int orc_search_by_bow_kf_f(const orc_bowframe* kf, const orc_bowframe* F, float nnratio, int check_ori, int32_t* match_f)
{
    ivec hist[HISTO_LENGTH]; memset(hist, 0, sizeof(hist));
    int nmatches = 0;
    for (int j = 0; j < F->n; ++j) match_f[j] = -1;
    FOR_SHARED_NODES(kf, F, ia, ib) {
        for (int a = kf->node_off[ia]; a < kf->node_off[ia + 1]; ++a) {
            const int iKF = kf->feat[a];
            if (!kf->valid[iKF]) continue;                                     /* :307-313 */
            const uint8_t* dKF = kf->desc + 32 * (size_t)iKF;
            int best1 = 256, best2 = 256, bestIdx = -1;
            for (int b = F->node_off[ib]; b < F->node_off[ib + 1]; ++b) {
                const int iF = F->feat[b];
                if (match_f[iF] >= 0) continue;                                /* :323 */
                const int dist = orc_descriptor_distance(dKF, F->desc + 32 * (size_t)iF);
                if (dist < best1) { best2 = best1; best1 = dist; bestIdx = iF; }
                else if (dist < best2) best2 = dist;
            }
            if (best1 <= TH_LOW && (float)best1 < nnratio * (float)best2) {    /* :348-351 */
                match_f[bestIdx] = iKF;
                if (check_ori) ipush(&hist[rot_bin(kf->keys[iKF].angle, F->keys[bestIdx].angle)], bestIdx);
                ++nmatches;
            }
        }
        ++ia; ++ib;
    }
    if (check_ori) {
        int cnt[HISTO_LENGTH], i1, i2, i3;
        for (int i = 0; i < HISTO_LENGTH; ++i) cnt[i] = hist[i].n;
        orc_three_maxima(cnt, HISTO_LENGTH, &i1, &i2, &i3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == i1 || i == i2 || i == i3) continue;
            for (int j = 0; j < hist[i].n; ++j) { match_f[hist[i].v[j]] = -1; --nmatches; }
        }
    }
    for (int i = 0; i < HISTO_LENGTH; ++i) free(hist[i].v);
    return nmatches;
}

/* SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12): src/ORBmatcher.cc:760-903.
 * match12[KF1.N] receives the KF2 feature index (whose MapPoint is stored in vpMatches12[idx1]) or -1. */
int orc_search_by_bow_kf_kf(const orc_bowframe* k1, const orc_bowframe* k2, float nnratio, int check_ori, int32_t* match12)
{
    ivec hist[HISTO_LENGTH]; memset(hist, 0, sizeof(hist));
    uint8_t* matched2 = (uint8_t*)calloc((size_t)(k2->n > 0 ? k2->n : 1), 1);
    int nmatches = 0;
    for (int j = 0; j < k1->n; ++j) match12[j] = -1;
    FOR_SHARED_NODES(k1, k2, ia, ib) {
        for (int a = k1->node_off[ia]; a < k1->node_off[ia + 1]; ++a) {
            const int idx1 = k1->feat[a];
            if (!k1->valid[idx1]) continue;
            const uint8_t* d1 = k1->desc + 32 * (size_t)idx1;
            int best1 = 256, best2 = 256, bestIdx2 = -1;
            for (int b = k2->node_off[ib]; b < k2->node_off[ib + 1]; ++b) {
                const int idx2 = k2->feat[b];
                if (matched2[idx2] || !k2->valid[idx2]) continue;              /* :822-826 */
                const int dist = orc_descriptor_distance(d1, k2->desc + 32 * (size_t)idx2);
                if (dist < best1) { best2 = best1; best1 = dist; bestIdx2 = idx2; }
                else if (dist < best2) best2 = dist;
            }
            if (best1 < TH_LOW && (float)best1 < nnratio * (float)best2) {     /* :845-847 (strict) */
                match12[idx1] = bestIdx2;
                matched2[bestIdx2] = 1;
                if (check_ori) ipush(&hist[rot_bin(k1->keys[idx1].angle, k2->keys[bestIdx2].angle)], idx1);
                ++nmatches;
            }
        }
        ++ia; ++ib;
    }
    if (check_ori) {
        int cnt[HISTO_LENGTH], i1, i2, i3;
        for (int i = 0; i < HISTO_LENGTH; ++i) cnt[i] = hist[i].n;
        orc_three_maxima(cnt, HISTO_LENGTH, &i1, &i2, &i3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == i1 || i == i2 || i == i3) continue;
            for (int j = 0; j < hist[i].n; ++j) { match12[hist[i].v[j]] = -1; --nmatches; }
        }
    }
    for (int i = 0; i < HISTO_LENGTH; ++i) free(hist[i].v);
    free(matched2);
    return nmatches;
}

/* CheckDistEpipolarLine: src/ORBmatcher.cc:213-240.  F12 row-major 3x3 float. */
static int check_epipolar(const orc_keypoint* kp1, const orc_keypoint* kp2, const float* F12, const float* sigma2)
{
    const float a = kp1->x * F12[0] + kp1->y * F12[3] + F12[6];
    const float b = kp1->x * F12[1] + kp1->y * F12[4] + F12[7];
    const float c = kp1->x * F12[2] + kp1->y * F12[5] + F12[8];
    const float num = a * kp2->x + b * kp2->y + c;
    const float den = a * a + b * b;
    if (den == 0) return 0;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * sigma2[kp2->octave];
}

/* SearchForTriangulation: src/ORBmatcher.cc:915-1127.  valid[i] = feature i already HAS a MapPoint
 * (those are skipped, :981,:1011).  (ex,ey) = epipole of KF1's centre in KF2 (:927-929, computed by
 * the caller).  match12[KF1.N] = KF2 index or -1 (vMatches12); pairs are its non-negative entries. */
int orc_search_for_triangulation(const orc_bowframe* k1, const orc_bowframe* k2, const float* F12, float ex, float ey,
                                 const float* scale2, const float* sigma2_2, int only_stereo, int check_ori, int32_t* match12)
{
    ivec hist[HISTO_LENGTH]; memset(hist, 0, sizeof(hist));
    uint8_t* matched2 = (uint8_t*)calloc((size_t)(k2->n > 0 ? k2->n : 1), 1);
    int nmatches = 0;
    for (int j = 0; j < k1->n; ++j) match12[j] = -1;
    FOR_SHARED_NODES(k1, k2, ia, ib) {
        for (int a = k1->node_off[ia]; a < k1->node_off[ia + 1]; ++a) {
            const int idx1 = k1->feat[a];
            if (k1->valid[idx1]) continue;                                     /* :981 */
            const int stereo1 = k1->uright && k1->uright[idx1] >= 0;
            if (only_stereo && !stereo1) continue;
            const orc_keypoint* kp1 = &k1->keys[idx1];
            const uint8_t* d1 = k1->desc + 32 * (size_t)idx1;
            int bestDist = TH_LOW, bestIdx2 = -1;
            for (int b = k2->node_off[ib]; b < k2->node_off[ib + 1]; ++b) {
                const int idx2 = k2->feat[b];
                if (matched2[idx2] || k2->valid[idx2]) continue;               /* :1011 */
                const int stereo2 = k2->uright && k2->uright[idx2] >= 0;
                if (only_stereo && !stereo2) continue;
                const int dist = orc_descriptor_distance(d1, k2->desc + 32 * (size_t)idx2);
                if (dist > TH_LOW || dist > bestDist) continue;                /* :1026 */
                const orc_keypoint* kp2 = &k2->keys[idx2];
                if (!stereo1 && !stereo2) {                                    /* :1034-1044 */
                    const float distex = ex - kp2->x, distey = ey - kp2->y;
                    if (distex * distex + distey * distey < 100 * scale2[kp2->octave]) continue;
                }
                if (check_epipolar(kp1, kp2, F12, sigma2_2)) { bestIdx2 = idx2; bestDist = dist; }
            }
            if (bestIdx2 >= 0) {
                match12[idx1] = bestIdx2;
                matched2[bestIdx2] = 1;                                        /* fork-specific, :1061 */
                ++nmatches;
                if (check_ori) ipush(&hist[rot_bin(kp1->angle, k2->keys[bestIdx2].angle)], idx1);
            }
        }
        ++ia; ++ib;
    }
    if (check_ori) {
        int cnt[HISTO_LENGTH], i1, i2, i3;
        for (int i = 0; i < HISTO_LENGTH; ++i) cnt[i] = hist[i].n;
        orc_three_maxima(cnt, HISTO_LENGTH, &i1, &i2, &i3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == i1 || i == i2 || i == i3) continue;
            for (int j = 0; j < hist[i].n; ++j) { match12[hist[i].v[j]] = -1; --nmatches; } /* :1108 also clears vbMatched2 */
        }
    }
    for (int i = 0; i < HISTO_LENGTH; ++i) free(hist[i].v);
    free(matched2);
    return nmatches;
}

/* ---------------- Frame grid ---------------- */
/* Frame::AssignFeaturesToGrid + PosInGrid: src/Frame.cc:398-422, 682-698.  CSR over cells, cell = ix*ROWS+iy. */
void orc_assign_grid(const orc_keypoint* keys, int n, const orc_grid* g, int32_t* cell_off, int32_t* cell_idx)
{
    const int nc = ORC_GRID_COLS * ORC_GRID_ROWS;
    int* cell = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    memset(cell_off, 0, sizeof(int32_t) * (size_t)(nc + 1));
    for (int i = 0; i < n; ++i) {
        const int px = (int)roundf((keys[i].x - g->min_x) * g->inv_w);
        const int py = (int)roundf((keys[i].y - g->min_y) * g->inv_h);
        cell[i] = (px < 0 || px >= ORC_GRID_COLS || py < 0 || py >= ORC_GRID_ROWS) ? -1 : px * ORC_GRID_ROWS + py;
        if (cell[i] >= 0) cell_off[cell[i] + 1]++;
    }
    for (int c = 0; c < nc; ++c) cell_off[c + 1] += cell_off[c];
    int* fill = (int*)calloc((size_t)nc, sizeof(int));
    for (int i = 0; i < n; ++i) if (cell[i] >= 0) cell_idx[cell_off[cell[i]] + fill[cell[i]]++] = i;
    free(fill); free(cell);
}

/* Frame::GetFeaturesInArea: src/Frame.cc:590-671 (KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:794-838, is the
 * same without the level filter: pass min_level = -1, max_level = -1). Returns count; out in visit order. */
int orc_features_in_area(const orc_keypoint* keys, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx,
                         float x, float y, float r, int min_level, int max_level, int32_t* out)
{
    int n = 0;
    int c0 = (int)floorf((x - g->min_x - r) * g->inv_w); if (c0 < 0) c0 = 0;
    if (c0 >= ORC_GRID_COLS) return 0;
    int c1 = (int)ceilf((x - g->min_x + r) * g->inv_w); if (c1 > ORC_GRID_COLS - 1) c1 = ORC_GRID_COLS - 1;
    if (c1 < 0) return 0;
    int r0 = (int)floorf((y - g->min_y - r) * g->inv_h); if (r0 < 0) r0 = 0;
    if (r0 >= ORC_GRID_ROWS) return 0;
    int r1 = (int)ceilf((y - g->min_y + r) * g->inv_h); if (r1 > ORC_GRID_ROWS - 1) r1 = ORC_GRID_ROWS - 1;
    if (r1 < 0) return 0;
    const int check = (min_level > 0) || (max_level >= 0);
    for (int ix = c0; ix <= c1; ++ix)
        for (int iy = r0; iy <= r1; ++iy) {
            const int c = ix * ORC_GRID_ROWS + iy;
            for (int j = cell_off[c]; j < cell_off[c + 1]; ++j) {
                const orc_keypoint* kp = &keys[cell_idx[j]];
                if (check) {
                    if (kp->octave < min_level) continue;
                    if (max_level >= 0 && kp->octave > max_level) continue;
                }
                const float dx = kp->x - x, dy = kp->y - y;
                if (fabsf(dx) < r && fabsf(dy) < r) out[n++] = cell_idx[j];
            }
        }
    return n;
}

/* SearchByProjection(Frame&, vector<MapPoint*>&, th): src/ORBmatcher.cc:78-191 + RadiusByViewingCos :194-201.
 * pts: the visible, good map points in caller order (mbTrackInView && !isBad).  occupied[j] != 0: frame
 * keypoint j already holds a MapPoint with Observations() > 0 (:136-138); it is updated as matches are made: the
 * reference writes F.mvpMapPoints[bestIdx] = pMP, which hides keypoint bestIdx from the following points iff pMP is
 * observed.  transient (optional): transient[i] != 0 = point i has Observations() == 0 (the temporal points of
 * Tracking::UpdateLastFrame); its match is recorded and counted but a later point may take the keypoint again.
 * assign[j] = index of the LAST point matched to keypoint j, or -1. */
int orc_search_by_projection_points(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                    const float* uright, const float* scale, const orc_grid* g, const int32_t* cell_off,
                                    const int32_t* cell_idx, const uint8_t* occupied_in, float th, float nnratio, int32_t* assign,
                                    const uint8_t* transient)
{
    uint8_t* occ = (uint8_t*)malloc((size_t)(n > 0 ? n : 1));
    int32_t* cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int j = 0; j < n; ++j) { occ[j] = occupied_in ? occupied_in[j] : 0; assign[j] = -1; }
    int nmatches = 0;
    const int bFactor = th != 1.0;
    for (int i = 0; i < npts; ++i) {
        const orc_projpoint* p = &pts[i];
        const int lvl = p->level;
        float r = p->view_cos > 0.998 ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        const int nc = orc_features_in_area(keys, g, cell_off, cell_idx, p->u, p->v, r * scale[lvl], lvl - 1, lvl, cand);
        if (nc == 0) continue;
        int best = 256, bestLevel = -1, best2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int k = 0; k < nc; ++k) {
            const int idx = cand[k];
            if (occ[idx]) continue;
            if (uright && uright[idx] > 0) {
                const float er = fabsf(p->ur - uright[idx]);
                if (er > r * scale[lvl]) continue;
            }
            const int dist = orc_descriptor_distance(p->desc, desc + 32 * (size_t)idx);
            if (dist < best) { best2 = best; best = dist; bestLevel2 = bestLevel; bestLevel = keys[idx].octave; bestIdx = idx; }
            else if (dist < best2) { bestLevel2 = keys[idx].octave; best2 = dist; }
        }
        if (best <= TH_HIGH) {
            if (bestLevel == bestLevel2 && (float)best > nnratio * (float)best2) continue;
            assign[bestIdx] = i;                                   /* F.mvpMapPoints[bestIdx] = pMP, :185 */
            if (!transient || !transient[i]) occ[bestIdx] = 1;     /* hidden from later points only if pMP is observed, :136-138 */
            ++nmatches;
        }
    }
    free(occ); free(cand);
    return nmatches;
}

/* SearchByProjection(Frame& Cur, const Frame& Last, th, bMono): src/ORBmatcher.cc:1741-1910, after the
 * projection (:1784-1808, done by the caller): pts[i] = (u, v, ur = u - mbf*invzc, level = nLastOctave,
 * angle = LastFrame.mvKeysUn[i].angle, desc).  motion: 0 = neither, 1 = forward, 2 = backward (:1768-1769). */
int orc_search_by_projection_frame(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                   const float* uright, const float* scale, const orc_grid* g, const int32_t* cell_off,
                                   const int32_t* cell_idx, const uint8_t* occupied_in, float th, int motion, int check_ori,
                                   int32_t* assign, const uint8_t* transient)
{
    ivec hist[HISTO_LENGTH]; memset(hist, 0, sizeof(hist));
    uint8_t* occ = (uint8_t*)malloc((size_t)(n > 0 ? n : 1));
    int32_t* cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int j = 0; j < n; ++j) { occ[j] = occupied_in ? occupied_in[j] : 0; assign[j] = -1; }
    int nmatches = 0;
    for (int i = 0; i < npts; ++i) {
        const orc_projpoint* p = &pts[i];
        const int lo = p->level;
        const float radius = th * scale[lo];
        int nc;
        if (motion == 1) nc = orc_features_in_area(keys, g, cell_off, cell_idx, p->u, p->v, radius, lo, -1, cand);
        else if (motion == 2) nc = orc_features_in_area(keys, g, cell_off, cell_idx, p->u, p->v, radius, 0, lo, cand);
        else nc = orc_features_in_area(keys, g, cell_off, cell_idx, p->u, p->v, radius, lo - 1, lo + 1, cand);
        if (nc == 0) continue;
        int best = 256, bestIdx2 = -1;
        for (int k = 0; k < nc; ++k) {
            const int i2 = cand[k];
            if (occ[i2]) continue;
            if (uright && uright[i2] > 0) {
                const float er = fabsf(p->ur - uright[i2]);
                if (er > radius) continue;
            }
            const int dist = orc_descriptor_distance(p->desc, desc + 32 * (size_t)i2);
            if (dist < best) { best = dist; bestIdx2 = i2; }
        }
        if (best <= TH_HIGH) {
            assign[bestIdx2] = i;                                  /* CurrentFrame.mvpMapPoints[bestIdx2] = pMP, :1858 */
            if (!transient || !transient[i]) occ[bestIdx2] = 1;    /* :1840-1842: skipped later only if pMP is observed */
            ++nmatches;
            if (check_ori) ipush(&hist[rot_bin(p->angle, keys[bestIdx2].angle)], bestIdx2);   /* one entry per accept */
        }
    }
    if (check_ori) {
        int cnt[HISTO_LENGTH], i1, i2, i3;
        for (int i = 0; i < HISTO_LENGTH; ++i) cnt[i] = hist[i].n;
        orc_three_maxima(cnt, HISTO_LENGTH, &i1, &i2, &i3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == i1 || i == i2 || i == i3) continue;
            /* :1893-1903: every entry clears its keypoint (whoever holds it by now) and un-counts itself */
            for (int j = 0; j < hist[i].n; ++j) { assign[hist[i].v[j]] = -1; --nmatches; }
        }
    }
    for (int i = 0; i < HISTO_LENGTH; ++i) free(hist[i].v);
    free(occ); free(cand);
    return nmatches;
}

/* SearchByProjection(Frame& Cur, KeyFrame*, set& sAlreadyFound, th, ORBdist): src/ORBmatcher.cc:1922-2065 (relocalisation),
 * after the caller's projection and PredictScale (:1949-1978): pts[i] = (u, v, level = nPredictedLevel,
 * angle = pKF->mvKeysUn[i].angle, desc).  Window levels [level-1, level+1] (:1985), keypoints already holding a
 * MapPoint skipped (:2000), best only, accept <= orb_dist (:2015), rotation histogram on the frame index (:2033). */
int orc_search_by_projection_reloc(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                   const float* scale, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx,
                                   const uint8_t* occupied_in, float th, int orb_dist, int check_ori, int32_t* assign)
{
    ivec hist[HISTO_LENGTH]; memset(hist, 0, sizeof(hist));
    uint8_t* occ = (uint8_t*)malloc((size_t)(n > 0 ? n : 1));
    int32_t* cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int j = 0; j < n; ++j) { occ[j] = occupied_in ? occupied_in[j] : 0; assign[j] = -1; }
    int nmatches = 0;
    for (int i = 0; i < npts; ++i) {
        const orc_projpoint* p = &pts[i];
        const int lvl = p->level;
        const float radius = th * scale[lvl];
        const int nc = orc_features_in_area(keys, g, cell_off, cell_idx, p->u, p->v, radius, lvl - 1, lvl + 1, cand);
        if (nc == 0) continue;
        int best = 256, bestIdx2 = -1;
        for (int k = 0; k < nc; ++k) {
            const int i2 = cand[k];
            if (occ[i2]) continue;
            const int dist = orc_descriptor_distance(p->desc, desc + 32 * (size_t)i2);
            if (dist < best) { best = dist; bestIdx2 = i2; }
        }
        if (best <= orb_dist) {
            assign[bestIdx2] = i; occ[bestIdx2] = 1;
            ++nmatches;
            if (check_ori) ipush(&hist[rot_bin(p->angle, keys[bestIdx2].angle)], bestIdx2);
        }
    }
    if (check_ori) {
        int cnt[HISTO_LENGTH], i1, i2, i3;
        for (int i = 0; i < HISTO_LENGTH; ++i) cnt[i] = hist[i].n;
        orc_three_maxima(cnt, HISTO_LENGTH, &i1, &i2, &i3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == i1 || i == i2 || i == i3) continue;
            for (int j = 0; j < hist[i].n; ++j) { assign[hist[i].v[j]] = -1; --nmatches; }
        }
    }
    for (int i = 0; i < HISTO_LENGTH; ++i) free(hist[i].v);
    free(occ); free(cand);
    return nmatches;
}

/* SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th): src/ORBmatcher.cc:431-560 (loop closing), after the
 * caller's Sim3 projection, depth / viewing-angle gates and PredictScale (:470-512): pts[i] = (u, v, level, desc).
 * KeyFrame::GetFeaturesInArea has no level filter (src/KeyFrame.cc:794-838); keypoints with vpMatched set are skipped
 * (:531), levels outside [level-1, level] are skipped (:536), best only, accept <= TH_LOW (:552), no histogram. */
int orc_search_by_projection_sim3(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                  const float* scale, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx,
                                  const uint8_t* occupied_in, int th, int32_t* assign)
{
    uint8_t* occ = (uint8_t*)malloc((size_t)(n > 0 ? n : 1));
    int32_t* cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int j = 0; j < n; ++j) { occ[j] = occupied_in ? occupied_in[j] : 0; assign[j] = -1; }
    int nmatches = 0;
    for (int i = 0; i < npts; ++i) {
        const orc_projpoint* p = &pts[i];
        const int lvl = p->level;
        const float radius = th * scale[lvl];
        const int nc = orc_features_in_area(keys, g, cell_off, cell_idx, p->u, p->v, radius, -1, -1, cand);
        if (nc == 0) continue;
        int best = 256, bestIdx = -1;
        for (int k = 0; k < nc; ++k) {
            const int idx = cand[k];
            if (occ[idx]) continue;
            const int kpLevel = keys[idx].octave;
            if (kpLevel < lvl - 1 || kpLevel > lvl) continue;
            const int dist = orc_descriptor_distance(p->desc, desc + 32 * (size_t)idx);
            if (dist < best) { best = dist; bestIdx = idx; }
        }
        if (best <= TH_LOW) { assign[bestIdx] = i; occ[bestIdx] = 1; ++nmatches; }
    }
    free(occ); free(cand);
    return nmatches;
}

/* SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize): src/ORBmatcher.cc:580-748.
 * keys1/desc1: F1.mvKeysUn / mDescriptors; prev_xy[2*i1..]: vbPrevMatched[i1]; the grid is F2's.  Only level-0 keypoints
 * of F1 search (:609), F2 candidates are level 0 inside the window (:615); a candidate whose current match is at least as
 * close is skipped (:643); best/second-best, accept best <= TH_LOW (:660) and best < second * nnratio (:664); a candidate
 * taken before is stolen (:667-671).  The histogram keeps the i1 of every accepted match, stolen or not (:692), exactly
 * like the reference, so the three maxima are computed on those counts.  matches12[i1] = i2 or -1; prev_xy is updated
 * for the matched i1 (:741-744). */
int orc_search_for_initialization(const orc_keypoint* keys1, const uint8_t* desc1, int n1, float* prev_xy,
                                  const orc_keypoint* keys2, const uint8_t* desc2, int n2, const orc_grid* g,
                                  const int32_t* cell_off, const int32_t* cell_idx, int window, float nnratio, int check_ori,
                                  int32_t* matches12)
{
    ivec hist[HISTO_LENGTH]; memset(hist, 0, sizeof(hist));
    int* matched_dist = (int*)malloc(sizeof(int) * (size_t)(n2 > 0 ? n2 : 1));
    int32_t* matches21 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n2 > 0 ? n2 : 1));
    int32_t* cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n2 > 0 ? n2 : 1));
    for (int j = 0; j < n2; ++j) { matched_dist[j] = INT_MAX; matches21[j] = -1; }
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    int nmatches = 0;
    for (int i1 = 0; i1 < n1; ++i1) {
        const int level1 = keys1[i1].octave;
        if (level1 > 0) continue;
        const int nc = orc_features_in_area(keys2, g, cell_off, cell_idx, prev_xy[2 * i1], prev_xy[2 * i1 + 1], (float)window,
                                            level1, level1, cand);
        if (nc == 0) continue;
        int best = INT_MAX, best2 = INT_MAX, bestIdx2 = -1;
        for (int k = 0; k < nc; ++k) {
            const int i2 = cand[k];
            const int dist = orc_descriptor_distance(desc1 + 32 * (size_t)i1, desc2 + 32 * (size_t)i2);
            if (matched_dist[i2] <= dist) continue;
            if (dist < best) { best2 = best; best = dist; bestIdx2 = i2; }
            else if (dist < best2) best2 = dist;
        }
        if (best <= TH_LOW) {
            if ((float)best < (float)best2 * nnratio) {
                if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; --nmatches; }
                matches12[i1] = bestIdx2; matches21[bestIdx2] = i1; matched_dist[bestIdx2] = best;
                ++nmatches;
                if (check_ori) ipush(&hist[rot_bin(keys1[i1].angle, keys2[bestIdx2].angle)], i1);
            }
        }
    }
    if (check_ori) {
        int cnt[HISTO_LENGTH], a, b, c;
        for (int i = 0; i < HISTO_LENGTH; ++i) cnt[i] = hist[i].n;
        orc_three_maxima(cnt, HISTO_LENGTH, &a, &b, &c);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == a || i == b || i == c) continue;
            for (int j = 0; j < hist[i].n; ++j) {
                const int idx1 = hist[i].v[j];
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; --nmatches; }
            }
        }
    }
    for (int i1 = 0; i1 < n1; ++i1)
        if (matches12[i1] >= 0) { prev_xy[2 * i1] = keys2[matches12[i1]].x; prev_xy[2 * i1 + 1] = keys2[matches12[i1]].y; }
    for (int i = 0; i < HISTO_LENGTH; ++i) free(hist[i].v);
    free(matched_dist); free(matches21); free(cand);
    return nmatches;
}

/* Candidate loops of Fuse(KeyFrame*, vector<MapPoint*>&, th) src/ORBmatcher.cc:1139-1300 (inv_sigma2 != NULL: reprojection gate
 * :1254-1285), Fuse(KeyFrame*, Scw, ...) :1310-1458 and SearchBySim3 :1472-1723 (inv_sigma2 == NULL), after the caller's
 * projection: pts[i] = (u, v, ur, level = nPredictedLevel, desc); radius = th * scale[level].  Queries are independent.
 * best_dist[i] = smallest distance among the candidates (-1: none), best_idx[i] = its keypoint if <= th_dist else -1. */
void orc_window_best(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, const float* uright, int n,
                     const float* scale, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx, float th,
                     const float* inv_sigma2, int th_dist, int32_t* best_idx, int32_t* best_dist)
{
    int32_t* cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < npts; ++i) {
        const orc_projpoint* p = &pts[i];
        const int lvl = p->level;
        const float radius = th * scale[lvl];
        const float u = p->u, v = p->v, ur = p->ur;
        const int nc = orc_features_in_area(keys, g, cell_off, cell_idx, u, v, radius, -1, -1, cand);
        int best = INT_MAX, bestIdx = -1;
        for (int k = 0; k < nc; ++k) {
            const int idx = cand[k];
            const int kpLevel = keys[idx].octave;
            if (kpLevel < lvl - 1 || kpLevel > lvl) continue;
            if (inv_sigma2) {
                if (uright && uright[idx] >= 0) {
                    const float ex = u - keys[idx].x, ey = v - keys[idx].y, er = ur - uright[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * inv_sigma2[kpLevel] > 7.8) continue;
                } else {
                    const float ex = u - keys[idx].x, ey = v - keys[idx].y;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * inv_sigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = orc_descriptor_distance(p->desc, desc + 32 * (size_t)idx);
            if (dist < best) { best = dist; bestIdx = idx; }
        }
        best_dist[i] = bestIdx >= 0 ? best : -1;
        best_idx[i] = (bestIdx >= 0 && best <= th_dist) ? bestIdx : -1;
    }
    free(cand);
}

/* MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:366-448: among the N observed descriptors of a map point, the
 * one whose median distance to all of them (itself included, index (int)(0.5 * (N - 1)) of the sorted row) is smallest;
 * the first such row wins.  desc: the points' descriptors back to back, point p owns rows [seg[p], seg[p+1]). */
static int cmp_int(const void* a, const void* b) { return *(const int*)a - *(const int*)b; }
void orc_distinctive_descriptors(const uint8_t* desc, const int32_t* seg, int npoints, int32_t* best)
{
    for (int p = 0; p < npoints; ++p) {
        const int N = seg[p + 1] - seg[p];
        best[p] = -1;
        if (N <= 0) continue;
        const uint8_t* D = desc + 32 * (size_t)seg[p];
        int* row = (int*)malloc(sizeof(int) * (size_t)N);
        int bestMedian = INT_MAX, bestIdx = 0;
        for (int i = 0; i < N; ++i) {
            for (int j = 0; j < N; ++j) row[j] = i == j ? 0 : orc_descriptor_distance(D + 32 * (size_t)i, D + 32 * (size_t)j);
            qsort(row, (size_t)N, sizeof(int), cmp_int);
            const int median = row[(int)(0.5 * (N - 1))];
            if (median < bestMedian) { bestMedian = median; bestIdx = i; }
        }
        best[p] = bestIdx;
        free(row);
    }
}

/* BowVector of TemplatedVocabulary::transform for a TF-IDF / L1 vocabulary (TemplatedVocabulary.h:1166-1221): every feature
 * whose word weight is > 0 does v.addWeight(id, w) in feature order (BowVector.cpp:37-49: insert, then +=), then
 * BowVector::normalize(L1) (BowVector.cpp:59-82: norm summed in ascending id, every value divided).  word_weight is indexed by
 * word id (WordValue = double).  Returns the number of words; ids ascending. */
int orc_bow_vector(const int32_t* word, int n, const double* word_weight, int nword_ids, int32_t* ids, double* values)
{
    double* acc = (double*)calloc((size_t)(nword_ids > 0 ? nword_ids : 1), sizeof(double));
    uint8_t* seen = (uint8_t*)calloc((size_t)(nword_ids > 0 ? nword_ids : 1), 1);
    for (int f = 0; f < n; ++f) {
        const int id = word[f];
        if (id < 0 || id >= nword_ids) continue;
        const double w = word_weight[id];
        if (!(w > 0)) continue;
        if (seen[id]) acc[id] += w; else { acc[id] = w; seen[id] = 1; }
    }
    int nw = 0;
    double norm = 0.0;
    for (int id = 0; id < nword_ids; ++id) if (seen[id]) { ids[nw] = id; values[nw] = acc[id]; norm += fabs(acc[id]); ++nw; }
    if (norm > 0.0) for (int i = 0; i < nw; ++i) values[i] /= norm;
    free(acc); free(seen);
    return nw;
}

/* L1Scoring::score, ScoringObject.cpp:28-61 */
double orc_bow_score_l1(const int32_t* ids1, const double* v1, int n1, const int32_t* ids2, const double* v2, int n2)
{
    int a = 0, b = 0;
    double score = 0;
    while (a < n1 && b < n2) {
        if (ids1[a] == ids2[b]) { const double vi = v1[a], wi = v2[b]; score += fabs(vi - wi) - fabs(vi) - fabs(wi); ++a; ++b; }
        else if (ids1[a] < ids2[b]) { while (a < n1 && ids1[a] < ids2[b]) ++a; }      /* lower_bound */
        else { while (b < n2 && ids2[b] < ids1[a]) ++b; }
    }
    return -score / 2.0;
}

/* ---------------- Frame::ComputeStereoMatches: src/Frame.cc:831-1128 ----------------
 * pyrL/pyrR: un-bordered level images packed back to back (level l at lvl_off[l], size lvl_w[l] x lvl_h[l]). */
typedef struct { int d, i; } distidx;
static int cmp_distidx(const void* a, const void* b)
{
    const distidx *x = (const distidx*)a, *y = (const distidx*)b;
    if (x->d != y->d) return x->d < y->d ? -1 : 1;
    return x->i < y->i ? -1 : (x->i > y->i);
}

int orc_stereo_matches(const orc_keypoint* kL, const uint8_t* dL, int N, const orc_keypoint* kR, const uint8_t* dR, int Nr,
                       const uint8_t* pyrL, const uint8_t* pyrR, const int* lvl_w, const int* lvl_h, const size_t* lvl_off,
                       const float* scale, const float* inv_scale, float mbf, float mb, float* uRight, float* depth)
{
    const int nRows = lvl_h[0];
    ivec* rows = (ivec*)calloc((size_t)nRows, sizeof(ivec));
    for (int i = 0; i < N; ++i) { uRight[i] = -1.0f; depth[i] = -1.0f; }
    for (int iR = 0; iR < Nr; ++iR) {                                          /* :869-888 */
        const float kpY = kR[iR].y;
        const float r = 2.0f * scale[kR[iR].octave];
        const int maxr = (int)ceilf(kpY + r), minr = (int)floorf(kpY - r);
        for (int yi = minr; yi <= maxr; ++yi) if (yi >= 0 && yi < nRows) ipush(&rows[yi], iR);
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    distidx* vd = (distidx*)malloc(sizeof(distidx) * (size_t)(N > 0 ? N : 1));
    int nvd = 0;
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    for (int iL = 0; iL < N; ++iL) {
        const orc_keypoint* kpL = &kL[iL];
        const int levelL = kpL->octave;
        const float vL = kpL->y, uL = kpL->x;
        const ivec* cands = &rows[(size_t)vL];
        if (cands->n == 0) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH; int bestIdxR = 0;
        for (int c = 0; c < cands->n; ++c) {
            const int iR = cands->v[c];
            const orc_keypoint* kpR = &kR[iR];
            if (kpR->octave < levelL - 1 || kpR->octave > levelL + 1) continue;
            const float uR = kpR->x;
            if (uR >= minU && uR <= maxU) {
                const int dist = orc_descriptor_distance(dL + 32 * (size_t)iL, dR + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = kR[bestIdxR].x;
            const float sf = inv_scale[kpL->octave];
            const float scaleduL = roundf(kpL->x * sf), scaledvL = roundf(kpL->y * sf), scaleduR0 = roundf(uR0 * sf);
            const int w = 5, L = 5;
            const int lw = lvl_w[kpL->octave];
            const uint8_t* IL = pyrL + lvl_off[kpL->octave];
            const uint8_t* IR = pyrR + lvl_off[kpL->octave];
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
            int best = INT_MAX, bestinc = 0;
            float vDists[11];
            const float iniu = scaleduR0 - L - w, endu = scaleduR0 + L + w + 1;   /* fork: - L - w, :1025 */
            if (iniu < 0 || endu >= (float)lw) continue;
            const float cL = (float)IL[(size_t)(y0 + w) * lw + xl0 + w];
            for (int inc = -L; inc <= L; ++inc) {
                const int xr0 = (int)(scaleduR0 + (float)inc - w);
                const float cR = (float)IR[(size_t)(y0 + w) * lw + xr0 + w];
                double s = 0;                                                   /* cv::norm L1 accumulates in double */
                for (int yy = 0; yy < 2 * w + 1; ++yy)
                    for (int xx = 0; xx < 2 * w + 1; ++xx) {
                        const float a = (float)IL[(size_t)(y0 + yy) * lw + xl0 + xx] - cL;
                        const float b = (float)IR[(size_t)(y0 + yy) * lw + xr0 + xx] - cR;
                        s += fabs((double)(a - b));
                    }
                const float dist = (float)s;
                if (dist < (float)best) { best = (int)dist; bestinc = inc; }
                vDists[L + inc] = dist;
            }
            if (bestinc == -L || bestinc == L) continue;
            const float d1 = vDists[L + bestinc - 1], d2 = vDists[L + bestinc], d3 = vDists[L + bestinc + 1];
            const float deltaR = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = scale[kpL->octave] * ((float)scaleduR0 + (float)bestinc + deltaR);
            float disparity = uL - bestuR;
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                vd[nvd].d = best; vd[nvd].i = iL; ++nvd;
            }
        }
    }
    if (nvd > 0) {                                                             /* :1112-1127 (reference: UB when empty) */
        qsort(vd, (size_t)nvd, sizeof(distidx), cmp_distidx);
        const float median = (float)vd[nvd / 2].d;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = nvd - 1; i >= 0; --i) {
            if ((float)vd[i].d < thDist) break;
            uRight[vd[i].i] = -1; depth[vd[i].i] = -1;
        }
    }
    for (int i = 0; i < nRows; ++i) free(rows[i].v);
    free(rows); free(vd);
    return nvd;
}
