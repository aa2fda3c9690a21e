/* orbx -- B200-native ORB front end: C ABI (the drop-in boundary).
 *
 * Everything the reference's ORBextractor / ORBmatcher hot path needs, as plain-C entry points
 * (pointers + sizes, no C++/torch types).  The reference-compatible C++ classes in
 * include/ORBextractor.h / include/ORBmatcher.h are thin wrappers over these.
 * Citations are relative to the reference checkout (QiuYue-bit/ORB-SLAM2-noted).
 *
 * Conventions: every call returns ORBX_OK (0) or a negative orbx_status; nothing throws or aborts.
 * The caller owns all in/out buffers; a ctx owns its device memory and one CUDA stream.  A ctx is
 * used by one thread at a time; distinct ctxs are independent (the reference runs two extractor
 * instances concurrently for stereo, src/Frame.cc:144-153).  There is NO CPU fallback: without a
 * CUDA device orbx_create fails with ORBX_E_CUDA.
 */
#ifndef ORBX_H
#define ORBX_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbx_ctx orbx_ctx;

/* Binary layout of cv::KeyPoint (28 bytes), as filled by src/ORBextractor.cc:1141-1186,1642-1656 */
typedef struct { float x, y, size, angle, response; int32_t octave, class_id; } orbx_keypoint;

typedef enum {
    ORBX_OK = 0,
    ORBX_E_INVALID = -1,     /* bad argument */
    ORBX_E_CUDA = -2,        /* CUDA runtime error (see orbx_last_cuda_error) */
    ORBX_E_NOMEM = -3,
    ORBX_E_CAPACITY = -4,    /* caller buffer too small / batch or image larger than the ctx was created for */
    ORBX_E_UNSUPPORTED = -5  /* geometry the reference itself cannot process (e.g. width/height < 0.5) */
} orbx_status;

enum { ORBX_IN_DEVICE = 1, ORBX_OUT_DEVICE = 2,   /* flags: where image / result buffers live */
       ORBX_OUT_DENSE = 4 };                      /* orbx_extract_batch with both of the above: see there */

#define ORBX_MAX_LEVELS 16

/* ---- lifetime.  Replaces ORBextractor::ORBextractor, src/ORBextractor.cc:471-579 ---- */
int orbx_create(orbx_ctx** out, int device, int nfeatures, float scale_factor, int nlevels,
                int ini_th_fast, int min_th_fast, int max_w, int max_h, int max_batch);
/* A light context for the matcher entry points (orbx_hamming_*, orbx_match_bow*, orbx_search_*, orbx_window_best,
 * orbx_assign_grid, orbx_distinctive_descriptors, orbx_bow_*, orbx_vocab_*): a device and a stream, no extraction arenas.
 * The reference constructs an ORBmatcher on the stack at every call site (src/Tracking.cc:1356, src/LocalMapping.cc:...),
 * so the drop-in keeps one of these per thread.  Extraction calls on it return ORBX_E_CAPACITY. */
int orbx_create_matcher(orbx_ctx** out, int device);
void orbx_destroy(orbx_ctx* ctx);
const char* orbx_strerror(int status);
const char* orbx_last_cuda_error(const orbx_ctx* ctx);

/* ---- getters.  Replace GetLevels/GetScaleFactor(s)/GetInverseScaleFactors/GetScaleSigmaSquares/
 *      GetInverseScaleSigmaSquares, include/ORBextractor.h:115-149.  Arrays hold nlevels entries. ---- */
int orbx_get_levels(const orbx_ctx* ctx);
float orbx_get_scale_factor(const orbx_ctx* ctx);
int orbx_get_tables(const orbx_ctx* ctx, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                    int* features_per_level);
/* per-frame cap that can never overflow: sum over levels of (quota + 3), rounded up to 32 */
int orbx_max_keypoints(const orbx_ctx* ctx);

/* ---- extraction.  Replaces ORBextractor::operator(), src/ORBextractor.cc:1543-1658
 *      (ComputePyramid :1664-1733, ComputeKeyPointsOctTree :1046-1195, DistributeOctTree :696-1042,
 *       IC_Angle :94-141, GaussianBlur :1617-1625, computeOrbDescriptor :153-204). ----
 * img: 8-bit single channel, `stride` bytes per row.  kps/desc receive n_out keypoints in the reference's
 * order (level 0..L-1, quadtree list order inside a level); desc is n x 32 bytes.  An empty image
 * (w or h == 0, or img == NULL) leaves outputs untouched and sets *n_out = 0, like :1547. */
int orbx_extract(orbx_ctx* ctx, const uint8_t* img, int w, int h, size_t stride, int flags,
                 orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out);
/* B frames of identical geometry, frame b at imgs + b*frame_stride.  Results for frame b are written at
 * kps + b*cap_per_frame, desc + b*cap_per_frame*32, n_out[b].  With ORBX_OUT_DEVICE the call is
 * asynchronous on the ctx stream (use orbx_sync). */
/* With ORBX_IN_DEVICE | ORBX_OUT_DEVICE | ORBX_OUT_DENSE the records are written DENSE: frame b's n_out[b] keypoints and
 * descriptors start at row sum(n_out[0..b)) of kps / desc instead of row b * cap_per_frame (the buffers are sized as before;
 * cap_per_frame still bounds a frame).  That is the layout orbx_gather_frames_dense sends without a compaction pass. */
int orbx_extract_batch(orbx_ctx* ctx, const uint8_t* imgs, int batch, int w, int h, size_t stride,
                       size_t frame_stride, int flags, orbx_keypoint* kps, uint8_t* desc,
                       int cap_per_frame, int* n_out);
/* Level image of frame `frame` of the LAST extract call (mvImagePyramid, include/ORBextractor.h:158).
 * border = 0: the w x h level; border = 19: with the reflect-101 frame of src/ORBextractor.cc:1704-1727.
 * dst is host memory unless flags has ORBX_OUT_DEVICE. */
int orbx_pyramid_level(orbx_ctx* ctx, int frame, int level, int border, int flags, uint8_t* dst,
                       size_t dst_stride, int* w, int* h);
/* All levels of frame `frame` of the LAST extract call at once, each with its reflect-101 frame of `border` pixels, packed
 * back to back: level l starts at offsets[l] (nlevels entries, bytes) with rows of w_l + 2 * border bytes -- the memory
 * behind ORBextractor::mvImagePyramid (src/ORBextractor.cc:1674-1727) in one kernel, one copy and one synchronisation
 * (orbx_pyramid_level costs one of each per level).  dst == NULL: sizing call, fills offsets[] and *total_bytes only. */
int orbx_pyramid_bordered(orbx_ctx* ctx, int frame, int border, int flags, uint8_t* dst, size_t dst_bytes, size_t* offsets,
                          size_t* total_bytes);
int orbx_level_size(const orbx_ctx* ctx, int level, int w0, int h0, int* w, int* h);

/* Colour to gray in front of the extractor.  Replaces the cvtColor calls of Tracking::GrabImageStereo / RGBD / Monocular
 * (src/Tracking.cc:252-278, 313-326, 377-390) with OpenCV's 8-bit fixed point (R*9798 + G*19235 + B*3735 + 16384) >> 15.
 * src: interleaved 3- or 4-channel 8-bit image; dst: w x h gray (a device dst can be passed to orbx_extract with
 * ORBX_IN_DEVICE). */
enum { ORBX_FMT_RGB = 0, ORBX_FMT_BGR = 1, ORBX_FMT_RGBA = 2, ORBX_FMT_BGRA = 3 };
int orbx_cvt_gray(orbx_ctx* ctx, const uint8_t* src, int w, int h, size_t stride, int format, int flags, uint8_t* dst,
                  size_t dst_stride);

/* Frame::UndistortKeyPoints, src/Frame.cc:725-781 (cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK): 5 iterations
 * in double precision, results rounded to float).  dist = mDistCoef (k1, k2, p1, p2[, k3]), ndist = 4 or 5; every other
 * field of a keypoint is copied.  dist[0] == 0 copies the keypoints unchanged, like :727-731. */
int orbx_undistort_keypoints(orbx_ctx* ctx, const orbx_keypoint* keys, int n, float fx, float fy, float cx, float cy,
                             const float* dist, int ndist, int flags, orbx_keypoint* out);

/* Frames per pipeline chunk of orbx_extract_batch when the images are HOST buffers (default 64, env ORBX_CHUNK; 0 = whole
 * batch in one pass): the H2D of chunk i+1 and the D2H of chunk i-1 overlap the kernels of chunk i.  Device-resident
 * batches always run in one pass. */
int orbx_set_chunk(orbx_ctx* ctx, int frames);

/* ---- stream / instrumentation (used by bench.py and tests; not part of the reference surface) ---- */
int orbx_sync(orbx_ctx* ctx);
void* orbx_stream(orbx_ctx* ctx);                 /* cudaStream_t */
int orbx_set_profiling(orbx_ctx* ctx, int on);     /* record CUDA events around every stage */
/* stage names: "pyramid","fast","octree","blur","describe","h2d","d2h","hamming", ... */
int orbx_stage_count(void);
const char* orbx_stage_name(int i);
int orbx_stage_stats(orbx_ctx* ctx, int i, double* total_ms, long long* launches); /* syncs; since last reset */
int orbx_reset_stats(orbx_ctx* ctx);
long long orbx_kernel_launches(const orbx_ctx* ctx); /* kernels launched by this ctx since creation */

/* ---- 256-bit Hamming matching.  Replaces ORBmatcher::DescriptorDistance src/ORBmatcher.cc:2123-2143
 *      and the best/second-best inner loops of the Search* functions. ---- */
/* Pairwise distances for n pairs (host or device pointers per flags). */
int orbx_descriptor_distance(orbx_ctx* ctx, const uint8_t* a, const uint8_t* b, int n, int flags, int32_t* dist);
/* Many-vs-many, segmented: query i in segment s is compared with every train j of the same segment
 * (segments = BoW nodes / row bands / the whole set when nseg == 1); q_seg/t_seg are CSR offsets
 * (nseg+1 entries) into q_idx/t_idx index lists (NULL lists = identity).  For each query row the
 * best (lowest distance, first train in list order on ties) and second-best distances are returned:
 * exactly the state of the loops at src/ORBmatcher.cc:325-345 before acceptance.
 * t_valid (optional, per train descriptor, 0 = skip) models the `continue`s at :323,:822-826. */
int orbx_hamming_best2(orbx_ctx* ctx, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt,
                       const int32_t* q_seg, const int32_t* q_idx, const int32_t* t_seg, const int32_t* t_idx,
                       int nseg, const uint8_t* t_valid, int flags,
                       int32_t* best_idx, int32_t* best_dist, int32_t* second_dist);
/* orbx_hamming_best2 for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same
 * arguments, same kernel, same results; it calls no cudaMalloc / cudaFree / allocating library routine on any path.
 * Its scratch (the CSR arrays, and the staging of host buffers when the flags say host) lives in one file-scope device
 * array of ORBX_HAMMING_STATIC_BYTES that every GPU loading the library holds until the process exits.
 * Limits: nseg <= ORBX_HAMMING_STATIC_MAX_SEG and a scratch need <= ORBX_HAMMING_STATIC_BYTES (3 x 256-byte-rounded
 * (nseg+1) x 4 B always; plus 32 B per descriptor, 4 B per index, 1 B per t_valid entry without ORBX_IN_DEVICE; plus 12 B
 * per query row without ORBX_OUT_DEVICE -- e.g. 8192 x 8192 host descriptors fit, any size fits with device buffers).
 * Beyond either limit it returns ORBX_E_CAPACITY and orbx_last_cuda_error() says which.  All calls on a GPU SHARE the
 * array: calls that overlap on the GPU (two contexts, streams or threads) race; serialise them.  The call returns after
 * its stream work has finished. */
#define ORBX_HAMMING_STATIC_BYTES (1 << 20)
#define ORBX_HAMMING_STATIC_MAX_SEG 1024
int orbx_hamming_best2_static(orbx_ctx* ctx, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt,
                              const int32_t* q_seg, const int32_t* q_idx, const int32_t* t_seg, const int32_t* t_idx,
                              int nseg, const uint8_t* t_valid, int flags,
                              int32_t* best_idx, int32_t* best_dist, int32_t* second_dist);

/* ---- DBoW2 vocabulary + transform (feeds SearchByBoW).  Replaces TemplatedVocabulary::transform,
 *      Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1145-1221,1255-1306 (FORB::distance, FORB.cpp:81-101). ----
 * The tree is passed flattened: node 0 = root; children of node i are child[child_off[i] .. child_off[i+1]) in file
 * order; a node without children is a leaf (word).  Vocabulary TRAINING (create/k-means) is out of scope. */
typedef struct orbx_vocab orbx_vocab;
int orbx_vocab_create(orbx_ctx* ctx, int k, int L, int nnodes, const int32_t* child_off, const int32_t* child,
                      const uint8_t* node_desc, const int32_t* word_id, const float* weight, orbx_vocab** out);
/* ORBvoc text format (TemplatedVocabulary::loadFromTextFile, TemplatedVocabulary.h:1385-1487) */
int orbx_vocab_load_text(orbx_ctx* ctx, const char* path, orbx_vocab** out);
/* scoring / weighting of the text header (DBoW2 ScoringType / WeightingType; 0 / 0 = L1_NORM / TF_IDF, what ORBvoc uses and
 * what orbx_vocab_create assumes).  orbx_bow_vector returns ORBX_E_UNSUPPORTED unless scoring is L1_NORM and weighting is
 * TF_IDF or TF: the other modes accumulate and normalise differently (TemplatedVocabulary.h:1166-1221). */
/* TemplatedVocabulary::saveToTextFile, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1492-1517: the vocabulary back into the
 * ORBvoc text format (byte for byte what the reference writes for the same tree). */
int orbx_vocab_save_text(const orbx_vocab* v, const char* path);
int orbx_vocab_modes(const orbx_vocab* v, int* scoring, int* weighting);
void orbx_vocab_destroy(orbx_vocab* v);
int orbx_vocab_info(const orbx_vocab* v, int* k, int* L, int* nnodes, int* nwords);
/* Per feature: leaf word id, leaf weight, and the node id at level L - levelsup (0 = root when levelsup >= L). */
int orbx_bow_transform(orbx_ctx* ctx, const orbx_vocab* v, const uint8_t* desc, int n, int levelsup, int flags,
                       int32_t* word, float* weight, int32_t* node);
/* FeatureVector (FeatureVector.cpp:37-54) as CSR from the per-feature node ids (features with weight <= 0 dropped,
 * TemplatedVocabulary.h:1184): ascending node id, ascending feature index inside a node.  Host index glue.
 * node_id[nn], node_off[nn+1], feat[<= n]; returns nn through *nnodes. */
/* The BowVector half of transform (TemplatedVocabulary.h:1166-1221: addWeight per feature, BowVector.cpp:37-49; then
 * BowVector::normalize(L1), BowVector.cpp:59-82) for TF-IDF / L1 vocabularies such as ORBvoc, and L1Scoring::score
 * (ScoringObject.cpp:28-61) of a query vector against many database vectors (KeyFrameDatabase::DetectLoopCandidates /
 * DetectRelocalizationCandidates, src/KeyFrameDatabase.cc).  WordValue is a double; sums run in the reference's order, so
 * results are bit-equal.  word = the word id of every feature (orbx_bow_transform, n <= 16384); ids / values (capacity n) receive the
 * vector in ascending word id, *nwords its size.  orbx_vocab_set_word_weights replaces the per-word weights by exact
 * doubles (orbx_vocab_create takes floats; the text loader calls it with the parsed values). */
int orbx_vocab_set_word_weights(orbx_vocab* v, const double* weights, int nwords);
int orbx_bow_vector(orbx_ctx* ctx, const orbx_vocab* v, const int32_t* word, int n, int flags, int32_t* ids, double* values,
                    int* nwords);
/* seg: HOST array of nvec + 1 offsets into ids / vals; score[j] = L1 score of the query against vector j. */
int orbx_bow_score_l1(orbx_ctx* ctx, const int32_t* q_ids, const double* q_vals, int nq, const int32_t* ids, const double* vals,
                      const int32_t* seg, int nvec, int flags, double* score);
int orbx_feature_vector(const int32_t* node, const float* weight, int n, int32_t* node_id, int32_t* node_off,
                        int32_t* feat, int* nnodes);

/* ---- greedy BoW matchers.  Replace ORBmatcher::SearchByBoW(KeyFrame*,Frame&,..) src/ORBmatcher.cc:255-419,
 *      SearchByBoW(KeyFrame*,KeyFrame*,..) :760-903 and SearchForTriangulation :915-1127 (CheckDistEpipolarLine
 *      :213-240, ComputeThreeMaxima :2076-2118), bit-equal including the order-dependent masking. ---- */
enum { ORBX_BOW_KF_FRAME = 0, ORBX_BOW_KF_KF = 1, ORBX_BOW_TRIANGULATE = 2 };
typedef struct {          /* one side of a match job; all pointers host or device per `flags` */
    const uint8_t* desc;          /* n x 32 */
    const orbx_keypoint* keys;    /* mvKeysUn: angle (all modes); x, y, octave (triangulation) */
    const uint8_t* has_point;     /* per feature: holds a good MapPoint (NULL = none do) */
    const float* uright;          /* mvuRight, or NULL for monocular */
    int n;
    int nnodes;                   /* FeatureVector CSR (orbx_feature_vector layout) */
    const int32_t* node_id; const int32_t* node_off; const int32_t* feat;   /* HOST arrays; a feat[] entry outside [0, n) -> ORBX_E_INVALID */
} orbx_bow_side;
typedef struct {          /* SearchForTriangulation geometry, computed by the caller as at :920-929 */
    float F12[9];                 /* row-major */
    float ex, ey;                 /* epipole of KF1's centre in KF2 */
    float scale2[ORBX_MAX_LEVELS], sigma2_2[ORBX_MAX_LEVELS]; /* pKF2->mvScaleFactors / mvLevelSigma2 */
    int only_stereo;
} orbx_triangulation;
/* match: KF_FRAME -> indexed by side-2 (Frame) feature, value = side-1 (KeyFrame) feature whose MapPoint was
 * assigned (vpMapPointMatches), -1 = none.  KF_KF / TRIANGULATE -> indexed by side-1 feature, value = side-2
 * feature (vpMatches12 / vMatches12), -1 = none.  Node lists are always HOST arrays. */
int orbx_match_bow(orbx_ctx* ctx, int mode, const orbx_bow_side* s1, const orbx_bow_side* s2,
                   const orbx_triangulation* tri, float nnratio, int check_orientation, int flags,
                   int32_t* match, int* nmatches);
/* The same for npairs independent keyframe pairs in one set of launches (the keyframe-pair jobs of LocalMapping /
 * LoopClosing): pair j uses s1[j], s2[j], tri[j] (tri may be NULL unless mode == TRIANGULATE) and writes
 * match[j] and nmatches[j].  Results are identical to npairs separate orbx_match_bow calls. */
int orbx_match_bow_batch(orbx_ctx* ctx, int mode, int npairs, const orbx_bow_side* const* s1, const orbx_bow_side* const* s2,
                         const orbx_triangulation* const* tri, float nnratio, int check_orientation, int flags,
                         int32_t* const* match, int* nmatches);

/* ---- window (projection) matchers.  Replace the four ORBmatcher::SearchByProjection overloads
 *      (src/ORBmatcher.cc:78-191, :431-560, :1741-1910, :1922-2065) after the caller's projection step,
 *      Frame::GetFeaturesInArea src/Frame.cc:590-671 / KeyFrame::GetFeaturesInArea src/KeyFrame.cc:794-838 and
 *      Frame::AssignFeaturesToGrid src/Frame.cc:398-422,682-698. ---- */
typedef struct {
    float u, v;            /* projected pixel */
    float radius;          /* search half-window (already multiplied by th and the level scale) */
    float ur;              /* projected right coordinate (used when the frame has uright > 0) */
    float angle;           /* keypoint angle of the source observation (rotation histogram) */
    int32_t min_level, max_level; /* octave filter, max_level < 0 = unbounded (GetFeaturesInArea semantics) */
    int32_t flags;         /* ORBX_PROJ_QUERY_*; 0 = the match hides its keypoint from the following queries */
    uint8_t desc[32];
} orbx_proj_query;
/* The query's map point has Observations() == 0 (the temporal points Tracking::UpdateLastFrame creates for stereo / RGB-D):
 * the reference skips a keypoint only while its CURRENT holder is observed (src/ORBmatcher.cc:136-138, :1840-1842), so the
 * match is recorded but later queries may take the same keypoint again: assign[] then holds the last one, nmatches counts
 * every accept, and with check_orientation every accept votes and un-counts itself when its bin is dropped (:1893-1903).
 * Only the map-point and motion-model overloads use it; relocalisation / Sim3 skip any held keypoint (:2000, :531). */
enum { ORBX_PROJ_QUERY_TRANSIENT = 1 };
typedef struct { float min_x, min_y, inv_w, inv_h; int cols, rows; } orbx_grid; /* mnMinX.. FRAME_GRID_COLS/ROWS */
enum { ORBX_PROJ_RATIO_SAME_LEVEL = 1,  /* keep best/second-best + the bestLevel==bestLevel2 ratio rule (:177-186) */
       ORBX_PROJ_CHECK_URIGHT = 2 };    /* |ur - uright[i]| > radius rejects (:141-151, :1846-1852) */
int orbx_assign_grid(orbx_ctx* ctx, const orbx_keypoint* keys, int n, const orbx_grid* g, int flags,
                     int32_t* cell_off /* cols*rows+1 */, int32_t* cell_idx /* n */);
/* Queries are processed in order; a keypoint taken by an earlier query (unless that query is ORBX_PROJ_QUERY_TRANSIENT)
 * or with occupied[i] != 0 on entry is skipped by later ones, exactly like the reference's writes to mvpMapPoints / vpMatched.
 * assign[i] = query index matched to keypoint i, or -1. */
int orbx_search_by_projection(orbx_ctx* ctx, const orbx_proj_query* q, int nq, const orbx_keypoint* keys,
                              const uint8_t* desc, const float* uright, int n, const orbx_grid* g,
                              const uint8_t* occupied, int mode_flags, int th_dist, float nnratio,
                              int check_orientation, int flags, int32_t* assign, int* nmatches);

/* ---- multi-GPU collection (no counterpart in the reference, which is one CPU process).  Frames and keyframe-pair jobs
 *      are sharded over one process per GPU with no exchange while they are computed; these calls are the FINAL GATHER of the
 *      keypoint / descriptor / match lists to one root rank over NCCL (NVLink).  Only the records in use travel: the counts are
 *      scanned and the padded blocks compacted on the device first.  NCCL is loaded at run time (libnccl.so.2): without it
 *      orbx_comm_create(nranks > 1) returns ORBX_E_UNSUPPORTED and everything else in this header still works.
 *      Set-up: rank 0 calls orbx_comm_unique_id and gets the 128 bytes to the other ranks by any means (MPI, a file,
 *      torch.distributed); every rank then calls orbx_comm_create with its context (the communicator lives on that context's
 *      device and has its own stream).  nranks == 1 needs no NCCL and no id. ---- */
typedef struct orbx_comm orbx_comm;
#define ORBX_COMM_ID_BYTES 128
int orbx_comm_unique_id(uint8_t* id /* ORBX_COMM_ID_BYTES */);
int orbx_comm_create(orbx_comm** out, orbx_ctx* ctx, int nranks, int rank, const uint8_t* id);
void orbx_comm_destroy(orbx_comm* comm);
const char* orbx_comm_last_error(const orbx_comm* comm);
/* Ordering.  orbx_comm_mark records "the results enqueued on ctx so far are complete" and returns a ticket (>= 0; at most 7
 * outstanding); a gather given that ticket starts after it, whatever the caller enqueues on ctx in between (the next step's
 * extraction into other buffers).  ticket < 0 = after everything enqueued on ctx at the time of the gather call.
 * orbx_comm_wait makes ctx's stream wait (on the device) for the gathers issued so far -- call it before overwriting buffers a
 * gather reads; orbx_comm_sync blocks the host until they are done. */
int orbx_comm_mark(orbx_comm* comm, orbx_ctx* ctx);
int orbx_comm_wait(orbx_comm* comm, orbx_ctx* ctx);
int orbx_comm_sync(orbx_comm* comm);
/* Frames: every rank passes the DEVICE buffers orbx_extract_batch(..., ORBX_OUT_DEVICE) filled (kps / desc: frames x cap rows,
 * n: frames counts).  At the root, all_n receives nranks x frames counts (rank-major) and all_kps / all_desc the records of rank 0's
 * frame 0, 1, ..., then rank 1's, ... back to back (device buffers of all_records_cap rows; ignored on the other ranks);
 * *total_records = rows written.  Every rank returns ORBX_E_CAPACITY when the ROOT's buffers are too small (nothing is sent;
 * all_records_cap of the other ranks is ignored, their all_* pointers may be NULL).
 * The call blocks the host once, for the ranks' totals; the transfer itself is asynchronous on the communicator's stream. */
int orbx_gather_frames(orbx_comm* comm, orbx_ctx* ctx, int ticket, const orbx_keypoint* kps, const uint8_t* desc, const int32_t* n,
                       int frames, int cap, int root, int32_t* all_n, orbx_keypoint* all_kps, uint8_t* all_desc,
                       size_t all_records_cap, long long* total_records);
/* The same for records that are dense already (orbx_extract_batch with ORBX_OUT_DENSE): no compaction pass, the blocks go
 * straight from kps / desc to the root (peer copy into its window, or ncclSend).  cap only clamps the counts. */
int orbx_gather_frames_dense(orbx_comm* comm, orbx_ctx* ctx, int ticket, const orbx_keypoint* kps, const uint8_t* desc, const int32_t* n,
                             int frames, int cap, int root, int32_t* all_n, orbx_keypoint* all_kps, uint8_t* all_desc,
                             size_t all_records_cap, long long* total_records);

/* Match lists: match = pairs x cap int32 (what orbx_match_bow_batch / orbx_search_by_projection write: row i matched to match[i],
 * -1 = none), rows[p] = rows in use of pair p; both DEVICE pointers.  At the root: all_n = nranks x pairs match counts and
 * all_pairs = the (row, match[row]) records (2 x int32 each, ascending row) of every pair in rank, then pair order. */
int orbx_gather_matches(orbx_comm* comm, orbx_ctx* ctx, int ticket, const int32_t* match, const int32_t* rows, int pairs, int cap,
                        int root, int32_t* all_n, int32_t* all_pairs, size_t all_records_cap, long long* total_records);

/* The peer window: device memory of this rank that the other ranks of the communicator map over NVLink / NVSwitch (CUDA IPC).
 * COLLECTIVE, once per communicator: every rank passes the bytes it wants to be able to receive as a root (0 = never a root)
 * and gets the base of its own window (*base, NULL for 0 bytes; owned by the communicator).  A gather whose root-side record
 * buffers (all_kps and all_desc, or all_pairs; all_n may lie anywhere) are carved out of the ROOT's window needs no staging
 * and no ncclSend/ncclRecv of records: each rank writes its dense records straight into the root's memory at their final place
 * (peer copies on the DMA engines; env ORBX_GATHER_COPY=0: the compaction kernel's own stores, ORBX_GATHER_CTAS of them), and
 * only the per-rank counts (the completion handshake) go through NCCL.  Results are the same either way.  All ranks return alike: ORBX_OK, or ORBX_E_UNSUPPORTED when some rank could not export or map a window (no peer
 * access, IPC not permitted) -- nothing stays mapped then and the gathers travel by ncclSend/ncclRecv as before.
 * orbx_comm_destroy unmaps and frees: call it on every rank once the last gather has completed EVERYWHERE (after a barrier of
 * the job), because a rank may not free its window while another still writes to it.  Env ORBX_GATHER_STAGED=1 at the root
 * turns the direct form off (for measurements).  No counterpart in the reference. */
int orbx_comm_window(orbx_comm* comm, size_t bytes, void** base);
/* How this communicator's gathers travelled so far: through the root's window / staged through ncclSend/ncclRecv. */
int orbx_comm_stats(const orbx_comm* comm, int* direct_gathers, int* staged_gathers);

/* ---- KeyFrameDatabase, src/KeyFrameDatabase.cc: the place-recognition database behind loop closing and relocalisation.
 *      Keyframes are named by the caller's ids (KeyFrame::mnId, 0 <= id < 2^28).  The reference keeps an inverted file of
 *      std::list<KeyFrame*>; here every keyframe's BowVector lies in one device arena and a query is a forward scan against
 *      a bitmap of the query's words (csrc/k_kfdb.cu) -- same candidates, same ORDER.  A database belongs to its ctx (device,
 *      stream) and is used by one thread at a time, like the reference's under its mutex.  Only vocabularies scored with
 *      L1_NORM (the ORB-SLAM2 vocabulary; see orbx_vocab_modes) are covered. ---- */
typedef struct orbx_kfdb orbx_kfdb;
int orbx_kfdb_create(orbx_ctx* ctx, int nwords, orbx_kfdb** out);            /* KeyFrameDatabase(voc) :38-42, nwords = voc.size() */
void orbx_kfdb_destroy(orbx_kfdb* db);
int orbx_kfdb_size(const orbx_kfdb* db);                                      /* keyframes currently held */
/* add(pKF) :50-56: words = the ascending word ids of pKF->mBowVec, vals their weights (HOST arrays of n entries) */
int orbx_kfdb_add(orbx_kfdb* db, int kf_id, const int32_t* words, const double* vals, int n);
int orbx_kfdb_erase(orbx_kfdb* db, int kf_id);                                /* erase(pKF) :64-83; unknown ids are ignored */
int orbx_kfdb_clear(orbx_kfdb* db);                                           /* clear() :85-89 */
/* What pKF->GetBestCovisibilityKeyFrames(10) returns (src/KeyFrame.cc:248-258): ids, strongest first (the first 10 are
 * kept).  Call it when the keyframe's ordered connections change; ids not in the database are fine (never met by a query). */
int orbx_kfdb_set_covisible(orbx_kfdb* db, int kf_id, const int32_t* ids, int n);
/* The same for n keyframes in one call: kf_ids[i]'s list is neigh[10 * i .. 10 * i + 10), padded with -1 (HOST arrays).  For a
 * caller without a hook in KeyFrame::UpdateBestCovisibles: refresh every keyframe's list in front of a query. */
int orbx_kfdb_set_covisible_bulk(orbx_kfdb* db, int n, const int32_t* kf_ids, const int32_t* neigh);
/* DetectLoopCandidates(pKF, minScore) :113-270: query_id = pKF->mnId, words / vals = pKF->mBowVec, connected = the ids of
 * pKF->GetConnectedKeyFrames().  out_ids = candidate ids in the reference's order; *n_out = how many there are (the first
 * cap are written).  The per-keyframe query state (mnLoopQuery, mnLoopWords, mLoopScore) persists between calls as in the
 * reference, including its behaviour for a query id used twice. */
int orbx_kfdb_detect_loop(orbx_kfdb* db, long long query_id, const int32_t* words, const double* vals, int nq, const int32_t* connected,
                          int nconnected, float min_score, int32_t* out_ids, int cap, int* n_out);
/* DetectRelocalizationCandidates(F) :275-399 for nqueries frames per call, in order (a later query sees the mRelocScore
 * values an earlier one left, :351).  query_ids[q] = F->mnId; query q's BowVector is words / vals [seg[q], seg[q+1]) (HOST
 * arrays; seg has nqueries + 1 entries, seg[0] = 0); its candidates go to out_ids[q * cap ...], their number to n_out[q].
 * One synchronisation per call.  mLoopScore / mRelocScore start at 0 (the reference leaves them uninitialised). */
int orbx_kfdb_detect_reloc(orbx_kfdb* db, int nqueries, const long long* query_ids, const int32_t* words, const double* vals,
                           const int32_t* seg, int32_t* out_ids, int cap, int* n_out);

/* ---- MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:366-448, for npoints map points per call: the observed
 *      descriptors of point p are rows [seg[p], seg[p+1]) of desc (seg: HOST array of npoints+1 offsets, at most 1024 rows
 *      per point); best[p] = row (relative to seg[p]) with the smallest median distance to the point's descriptors, first
 *      one on ties; -1 for a point without descriptors. ---- */
int orbx_distinctive_descriptors(orbx_ctx* ctx, const uint8_t* desc, const int32_t* seg, int npoints, int flags, int32_t* best);

/* ---- independent window searches: every query finds its best keypoint on its own (no claims between queries).  Replaces
 *      the candidate loops of ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>&, th) src/ORBmatcher.cc:1139-1300 (with the
 *      reprojection gate :1254-1285: inv_sigma2 = pKF->mvInvLevelSigma2, a HOST array of ORBX_MAX_LEVELS floats; a keypoint
 *      with uright >= 0 uses the stereo form and 7.8, otherwise 5.99), Fuse(KeyFrame*, Scw, ..) :1310-1458 and both
 *      directions of SearchBySim3 :1472-1723 (inv_sigma2 = NULL: no gate).  The caller projects (u, v, ur, radius,
 *      min_level = L-1, max_level = L).  best_dist[i] = smallest distance in the window or -1 when it is empty;
 *      best_idx[i] = its keypoint (first visited wins a tie) when best_dist[i] <= th_dist, else -1. ---- */
int orbx_window_best(orbx_ctx* ctx, const orbx_proj_query* q, int nq, const orbx_keypoint* keys, const uint8_t* desc,
                     const float* uright, int n, const orbx_grid* g, const float* inv_sigma2, int th_dist, int flags,
                     int32_t* best_idx, int32_t* best_dist);

/* ---- monocular initialisation.  Replaces ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:580-748 (window search of
 *      F1's level-0 keypoints around vbPrevMatched in F2's grid, best/second-best ratio test, distance-masked candidates,
 *      stolen matches, rotation histogram).  prev_xy = vbPrevMatched as n1 (x, y) pairs, updated in place for the matched
 *      keypoints (:741-744); matches12[i1] = F2 index or -1 (vnMatches12).  With ORBX_IN_DEVICE every input pointer
 *      (prev_xy included) is a device pointer; ORBX_OUT_DEVICE leaves matches12 on the device. ---- */
int orbx_search_for_initialization(orbx_ctx* ctx, const orbx_keypoint* keys1, const uint8_t* desc1, int n1, float* prev_xy,
                                   const orbx_keypoint* keys2, const uint8_t* desc2, int n2, const orbx_grid* g,
                                   int window, float nnratio, int check_orientation, int flags, int32_t* matches12,
                                   int* nmatches);

/* ---- stereo.  Replaces Frame::ComputeStereoMatches, src/Frame.cc:831-1128.  left/right are the two extractor
 *      contexts whose LAST orbx_extract call produced the keypoints (their pyramids are read on the device). ---- */
int orbx_stereo_match(orbx_ctx* left, orbx_ctx* right, int frame, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                      int n_l, const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r, float bf, float b,
                      int flags, float* u_right, float* depth, int* n_matched);
/* n_matched (optional) = the left keypoints that END with a depth, i.e. after the 1.5 * 1.4 * median SAD rejection (:1112-1127).
 * Keypoints whose octave is outside [0, nlevels) or whose 11 x 11 SAD window leaves its level are left unmatched (the
 * reference's rowRange would assert); at most 2^20 - 1 keypoints per image (ORBX_E_CAPACITY). */
/* The same for `pairs` stereo pairs in one set of launches (configs[1] / configs[2] as throughput jobs): pair p uses frames
 * frame0 + p of the two contexts' last extract calls and rows [p * cap, p * cap + n[p]) of the arrays, the layout
 * orbx_extract_batch writes (cap = its cap_per_frame).  With ORBX_IN_DEVICE every input -- n_l and n_r included -- is a device
 * pointer, so it chains behind two orbx_extract_batch(..., ORBX_OUT_DEVICE) calls without a host round trip.  u_right / depth:
 * pairs x cap floats.  n_matched: host array of `pairs` counts or NULL (NULL keeps a device-to-device call asynchronous). */
int orbx_stereo_match_batch(orbx_ctx* left, orbx_ctx* right, int frame0, int pairs, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                            const int32_t* n_l, const orbx_keypoint* kps_r, const uint8_t* desc_r, const int32_t* n_r, int cap,
                            float bf, float b, int flags, float* u_right, float* depth, int32_t* n_matched);

#ifdef __cplusplus
}
#endif
#endif
