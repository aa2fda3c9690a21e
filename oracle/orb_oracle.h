/* TEST INFRASTRUCTURE ONLY (oracle/).  CPU restatement of the reference hot path:
 * ORBextractor::operator() and the Hamming core of ORBmatcher
 * (QiuYue-bit/ORB-SLAM2-noted; file:line citations are relative to /root/reference).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (liborbx.so) never links or calls it.
 *
 * PARITY PINNING: pinned.  (1) Control flow: tests/test_oracle_vs_ref.py compares this library byte for byte with the
 * reference's own sources compiled unmodified (oracle/_ref/liborb_ref.so, recipe in oracle/Makefile, cv:: shim in
 * oracle/cvshim/): whole frames, pyramids, the quadtree, stereo frames, all twelve ORBmatcher methods, DBoW2 transform.
 * (2) The third-party arithmetic the reference calls (cv::resize, cv::GaussianBlur, cv::FAST, cv::fastAtan2,
 * cv::copyMakeBorder, cv::undistortPoints, cv::cvtColor, libm cosf/sinf) against the executable copies in this image
 * (cv2 4.13.0, glibc 2.39): tests/golden fixtures generated with tools/gen_golden.py.  The reference has no tests of its
 * own for this path.  The quadtree tie-break is canonicalised (see orc_distribute_octree): the reference's depends on heap
 * addresses; it equals this library when address order == creation order, and DESIGN.md section 2 holds the measured
 * difference under glibc's real heap.
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_LEVELS 16
#define ORC_EDGE_THRESHOLD 19 /* src/ORBextractor.cc:81 */
#define ORC_HALF_PATCH 15     /* :80 */
#define ORC_PATCH 31          /* :79 */

typedef struct { float x, y, size, angle, response; int32_t octave, class_id; } orc_keypoint; /* == cv::KeyPoint, 28 B */

typedef struct {
    int nfeatures, nlevels, ini_th, min_th;
    float scale_factor;
    float scale[ORC_MAX_LEVELS], inv_scale[ORC_MAX_LEVELS];
    float sigma2[ORC_MAX_LEVELS], inv_sigma2[ORC_MAX_LEVELS];
    int nfeat_level[ORC_MAX_LEVELS];
    int umax[ORC_HALF_PATCH + 1];
    int blur_ksize; /* 9 in this fork (src/ORBextractor.cc:1622); 7 upstream */
} orc_params;

/* ctor tables, src/ORBextractor.cc:471-579 */
int orc_params_init(orc_params* p, int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
/* level sizes, src/ORBextractor.cc:1669-1672 */
void orc_level_size(const orc_params* p, int level, int w0, int h0, int* w, int* h);

/* primitives (third-party OpenCV arithmetic restated; pinned vs cv2 4.13 goldens) */
void orc_resize_linear(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh, size_t dstride);
void orc_border_reflect101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride, int border);
void orc_gauss_blur(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride, int ksize);
float orc_fast_atan2(float y, float x);
int orc_fast_score(const uint8_t* center, size_t stride, int threshold);
/* cv::FAST(roi, kps, th, nms=true) on a w x h ROI: writes (x,y,score) triplets, returns count */
int orc_fast_roi(const uint8_t* roi, int w, int h, size_t stride, int threshold, int* xys, int cap);
float orc_sinf(float x);
float orc_cosf(float x);

/* stages */
/* FAST over all cells of one level (src/ORBextractor.cc:1058-1153): window-relative x,y and response */
int orc_detect_level(const orc_params* p, const uint8_t* img, int w, int h, size_t stride, orc_keypoint* out, int cap);
/* DistributeOctTree :696-1042 (canonical tie-break). in: window-relative kps in detection order */
int orc_distribute_octree(const orc_keypoint* in, int n, int minX, int maxX, int minY, int maxY, int N, orc_keypoint* out, int cap);
float orc_ic_angle(const uint8_t* img, size_t stride, int cx, int cy, const int* umax); /* :94-141 */
void orc_brief(const uint8_t* blurred, size_t stride, int cx, int cy, float angle_deg, uint8_t* desc32); /* :153-204 */

/* full operator(): returns number of keypoints (<= cap), or <0 on error.
 * pyr (optional) receives the un-bordered levels packed back to back (sum w_l*h_l bytes). */
int orc_extract(const orc_params* p, const uint8_t* img, int w, int h, size_t stride,
                orc_keypoint* kps, uint8_t* desc, int cap, uint8_t* pyr);

/* ---- matcher side (orb_match.c) ---- */
void orc_cvt_gray(const uint8_t* src, int w, int h, size_t sstride, int channels, int bgr, uint8_t* dst, size_t dstride); /* cvtColor */
void orc_undistort_keypoints(const orc_keypoint* in, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                             orc_keypoint* out); /* src/Frame.cc:725-781 */
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b); /* src/ORBmatcher.cc:2123-2143 */
void orc_hamming_best2(const uint8_t* q, const uint8_t* t, const int32_t* q_seg, const int32_t* q_idx, const int32_t* t_seg,
                       const int32_t* t_idx, int nseg, const uint8_t* t_valid, int32_t* best_idx, int32_t* best_dist,
                       int32_t* second_dist); /* inner loops :325-345 */
void orc_three_maxima(const int* count, int L, int* ind1, int* ind2, int* ind3); /* :2076-2118 */

/* DBoW2 vocabulary tree flattened: node 0 = root; children of node i = child[child_off[i]..child_off[i+1])
 * in file order (TemplatedVocabulary.h:1385-1487); leaf <=> no children. */
typedef struct {
    int k, L, nnodes;
    const int32_t* child_off; const int32_t* child;
    const uint8_t* desc;       /* nnodes x 32 */
    const int32_t* word_id;    /* leaves only */
    const float* weight;
} orc_vocab;
void orc_bow_transform(const orc_vocab* V, const uint8_t* desc, int n, int levelsup, int32_t* word, float* weight, int32_t* node);
int orc_feature_vector(const int32_t* node, const float* weight, int n, int32_t* node_id, int32_t* node_off, int32_t* feat);

/* one (key)frame as the BoW matchers see it */
typedef struct {
    int n;
    const uint8_t* desc;        /* n x 32 */
    const orc_keypoint* keys;   /* mvKeysUn */
    const uint8_t* valid;       /* per feature: has a good MapPoint */
    const float* uright;        /* mvuRight or NULL (monocular) */
    int nnodes;                 /* FeatureVector in CSR, ascending node id */
    const int32_t* node_id; const int32_t* node_off; const int32_t* feat;
} orc_bowframe;
int orc_search_by_bow_kf_f(const orc_bowframe* kf, const orc_bowframe* F, float nnratio, int check_ori, int32_t* match_f);
int orc_search_by_bow_kf_kf(const orc_bowframe* k1, const orc_bowframe* k2, float nnratio, int check_ori, int32_t* match12);
int orc_search_for_triangulation(const orc_bowframe* k1, const orc_bowframe* k2, const float* F12, float ex, float ey,
                                 const float* scale2, const float* sigma2_2, int only_stereo, int check_ori, int32_t* match12);

#define ORC_GRID_COLS 64 /* include/Frame.h:55 */
#define ORC_GRID_ROWS 48 /* include/Frame.h:60 */
typedef struct { float min_x, min_y, inv_w, inv_h; } orc_grid; /* mnMinX, mnMinY, mfGridElementWidthInv/HeightInv */
void orc_assign_grid(const orc_keypoint* keys, int n, const orc_grid* g, int32_t* cell_off, int32_t* cell_idx);
int orc_features_in_area(const orc_keypoint* keys, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx,
                         float x, float y, float r, int min_level, int max_level, int32_t* out);
/* a projected map point (MapPoint::mTrackProjX/Y/XR, mnTrackScaleLevel, mTrackViewCos, GetDescriptor) */
typedef struct { float u, v, ur, view_cos, angle; int32_t level; uint8_t desc[32]; } orc_projpoint;
int orc_search_by_projection_points(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                    const float* uright, const float* scale, const orc_grid* g, const int32_t* cell_off,
                                    const int32_t* cell_idx, const uint8_t* occupied_in, float th, float nnratio, int32_t* assign,
                                    const uint8_t* transient /* optional: point i has Observations() == 0 */);
int orc_search_by_projection_frame(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                   const float* uright, const float* scale, const orc_grid* g, const int32_t* cell_off,
                                   const int32_t* cell_idx, const uint8_t* occupied_in, float th, int motion, int check_ori,
                                   int32_t* assign, const uint8_t* transient);
int orc_search_by_projection_reloc(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                   const float* scale, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx,
                                   const uint8_t* occupied_in, float th, int orb_dist, int check_ori, int32_t* assign); /* :1922-2065 */
int orc_search_by_projection_sim3(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, int n,
                                  const float* scale, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx,
                                  const uint8_t* occupied_in, int th, int32_t* assign); /* :431-560 */
int orc_search_for_initialization(const orc_keypoint* keys1, const uint8_t* desc1, int n1, float* prev_xy,
                                  const orc_keypoint* keys2, const uint8_t* desc2, int n2, const orc_grid* g,
                                  const int32_t* cell_off, const int32_t* cell_idx, int window, float nnratio, int check_ori,
                                  int32_t* matches12); /* :580-748 */
void orc_window_best(const orc_projpoint* pts, int npts, const orc_keypoint* keys, const uint8_t* desc, const float* uright, int n,
                     const float* scale, const orc_grid* g, const int32_t* cell_off, const int32_t* cell_idx, float th,
                     const float* inv_sigma2, int th_dist, int32_t* best_idx, int32_t* best_dist); /* Fuse x2, SearchBySim3 */
void orc_distinctive_descriptors(const uint8_t* desc, const int32_t* seg, int npoints, int32_t* best); /* src/MapPoint.cc:366-448 */
int orc_bow_vector(const int32_t* word, int n, const double* word_weight, int nword_ids, int32_t* ids, double* values); /* transform's BowVector */
double orc_bow_score_l1(const int32_t* ids1, const double* v1, int n1, const int32_t* ids2, const double* v2, int n2); /* ScoringObject.cpp:28-61 */
int orc_stereo_matches(const orc_keypoint* kL, const uint8_t* dL, int N, const orc_keypoint* kR, const uint8_t* dR, int Nr,
                       const uint8_t* pyrL, const uint8_t* pyrR, const int* lvl_w, const int* lvl_h, const size_t* lvl_off,
                       const float* scale, const float* inv_scale, float mbf, float mb, float* uRight, float* depth);

/* ---- KeyFrameDatabase (src/KeyFrameDatabase.cc): orb_kfdb.c ---- */
typedef struct orc_kfdb orc_kfdb;
orc_kfdb* orc_kfdb_create(int nwords);
void orc_kfdb_destroy(orc_kfdb* d);
void orc_kfdb_clear(orc_kfdb* d);
int orc_kfdb_new_keyframe(orc_kfdb* d, int id, const int32_t* word, const double* val, int n);
int orc_kfdb_add(orc_kfdb* d, int id);
int orc_kfdb_erase(orc_kfdb* d, int id);
int orc_kfdb_set_covisible(orc_kfdb* d, int id, const int32_t* ids, int n);
int orc_kfdb_detect_loop(orc_kfdb* d, int query_kf, const int32_t* conn, int nconn, float minScore, int32_t* out, int cap);
int orc_kfdb_detect_reloc(orc_kfdb* d, long long frame_id, const int32_t* word, const double* val, int n, int32_t* out, int cap);

#ifdef __cplusplus
}
#endif
#endif
