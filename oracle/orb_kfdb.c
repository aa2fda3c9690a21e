/* TEST INFRASTRUCTURE (the checker, never the product): CPU restatement of the reference's KeyFrameDatabase
 * (src/KeyFrameDatabase.cc) -- the inverted file over BoW words and the two candidate searches built on it.
 * Parity PINNED: tests/test_oracle_vs_ref.py runs the same scripts of add / erase / query through the unmodified
 * KeyFrameDatabase of oracle/_ref and through this file.
 *
 * The restatement keeps the reference's data flow literally: one posting list per word in insertion order
 * (mvInvertedFile, :50-56), the per-keyframe query state the reference stores IN the keyframes (mnLoopQuery / mnLoopWords /
 * mLoopScore, mnRelocQuery / mnRelocWords / mRelocScore, include/KeyFrame.h:322-333) and the std::list order of every
 * intermediate list, because the output is an ordered vector.  Two values the reference leaves uninitialised
 * (mLoopScore, mRelocScore: src/KeyFrame.cc:43-60 has no initialiser) start at 0 here; the _ref harness sets the same. */
#include "orb_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int exists, in_db;
    int nbow; int32_t* word; double* val;                 /* mBowVec: ascending word ids */
    int nneigh; int32_t neigh[10];                        /* GetBestCovisibilityKeyFrames(10): ids, strongest first */
    long long loop_query, reloc_query; int loop_words, reloc_words; float loop_score, reloc_score;
} OrcKf;

struct orc_kfdb {
    int nwords;
    int32_t** post; int* npost; int* cpost;               /* mvInvertedFile[word]: keyframe ids in insertion order */
    OrcKf* kf; int nkf;                                   /* indexed by keyframe id */
};

orc_kfdb* orc_kfdb_create(int nwords)                     /* KeyFrameDatabase::KeyFrameDatabase, :38-42 */
{
    orc_kfdb* d = (orc_kfdb*)calloc(1, sizeof(*d));
    d->nwords = nwords;
    d->post = (int32_t**)calloc((size_t)nwords, sizeof(int32_t*));
    d->npost = (int*)calloc((size_t)nwords, sizeof(int));
    d->cpost = (int*)calloc((size_t)nwords, sizeof(int));
    return d;
}

void orc_kfdb_clear(orc_kfdb* d)                          /* clear, :85-89 (the keyframes themselves live on) */
{
    for (int w = 0; w < d->nwords; ++w) d->npost[w] = 0;
    for (int i = 0; i < d->nkf; ++i) d->kf[i].in_db = 0;
}

void orc_kfdb_destroy(orc_kfdb* d)
{
    if (!d) return;
    for (int w = 0; w < d->nwords; ++w) free(d->post[w]);
    for (int i = 0; i < d->nkf; ++i) { free(d->kf[i].word); free(d->kf[i].val); }
    free(d->post); free(d->npost); free(d->cpost); free(d->kf); free(d);
}

static OrcKf* kf_slot(orc_kfdb* d, int id)
{
    if (id < 0) return NULL;
    if (id >= d->nkf) {
        int n = d->nkf ? d->nkf : 64;
        while (n <= id) n *= 2;
        d->kf = (OrcKf*)realloc(d->kf, sizeof(OrcKf) * (size_t)n);
        memset(d->kf + d->nkf, 0, sizeof(OrcKf) * (size_t)(n - d->nkf));
        d->nkf = n;
    }
    return &d->kf[id];
}

/* a keyframe object (it exists whether or not the database holds it) */
int orc_kfdb_new_keyframe(orc_kfdb* d, int id, const int32_t* word, const double* val, int n)
{
    OrcKf* k = kf_slot(d, id);
    if (!k || k->exists) return -1;
    for (int i = 0; i < n; ++i) if (word[i] < 0 || word[i] >= d->nwords || (i && word[i] <= word[i - 1])) return -1;
    memset(k, 0, sizeof(*k));
    k->exists = 1; k->nbow = n;
    k->word = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n ? n : 1)); k->val = (double*)malloc(sizeof(double) * (size_t)(n ? n : 1));
    memcpy(k->word, word, sizeof(int32_t) * (size_t)n); memcpy(k->val, val, sizeof(double) * (size_t)n);
    return 0;
}

int orc_kfdb_add(orc_kfdb* d, int id)                     /* add, :50-56 */
{
    OrcKf* k = kf_slot(d, id);
    if (!k || !k->exists) return -1;
    for (int i = 0; i < k->nbow; ++i) {
        const int w = k->word[i];
        if (d->npost[w] == d->cpost[w]) {
            d->cpost[w] = d->cpost[w] ? 2 * d->cpost[w] : 4;
            d->post[w] = (int32_t*)realloc(d->post[w], sizeof(int32_t) * (size_t)d->cpost[w]);
        }
        d->post[w][d->npost[w]++] = id;
    }
    k->in_db = 1;
    return 0;
}

int orc_kfdb_erase(orc_kfdb* d, int id)                   /* erase, :64-83: the first occurrence in every list of its words */
{
    OrcKf* k = kf_slot(d, id);
    if (!k || !k->exists) return -1;
    for (int i = 0; i < k->nbow; ++i) {
        const int w = k->word[i];
        for (int j = 0; j < d->npost[w]; ++j)
            if (d->post[w][j] == id) {
                memmove(d->post[w] + j, d->post[w] + j + 1, sizeof(int32_t) * (size_t)(d->npost[w] - j - 1));
                d->npost[w]--;
                break;
            }
    }
    k->in_db = 0;
    return 0;
}

int orc_kfdb_set_covisible(orc_kfdb* d, int id, const int32_t* ids, int n)   /* what GetBestCovisibilityKeyFrames(10) will return */
{
    OrcKf* k = kf_slot(d, id);
    if (!k || !k->exists) return -1;
    k->nneigh = n < 10 ? n : 10;
    for (int i = 0; i < k->nneigh; ++i) k->neigh[i] = ids[i];
    return 0;
}

/* L1Scoring::score, Thirdparty/DBoW2/DBoW2/ScoringObject.cpp:23-71 (the lower_bound jumps visit the same common words in
 * the same order as this merge) */
static double l1_score(const int32_t* w1, const double* v1, int n1, const int32_t* w2, const double* v2, int n2)
{
    int a = 0, b = 0;
    double s = 0;
    while (a < n1 && b < n2) {
        if (w1[a] == w2[b]) { s += fabs(v1[a] - v2[b]) - fabs(v1[a]) - fabs(v2[b]); ++a; ++b; }
        else if (w1[a] < w2[b]) ++a; else ++b;
    }
    return -s / 2.0;
}

typedef struct { float score; int id; } ScoreKf;

/* DetectLoopCandidates :113-270 (loop = 1: query is keyframe `qid` with its connected keyframes, minScore gate) and
 * DetectRelocalizationCandidates :275-399 (loop = 0: query is frame `qid`) */
static int detect(orc_kfdb* d, int loop, long long qid, const int32_t* qw, const double* qv, int nq, const int32_t* conn, int nconn,
                  float minScore, int32_t* out, int cap)
{
    int* sharing = (int*)malloc(sizeof(int) * (size_t)(d->nkf + 1));
    int ns = 0;
    for (int i = 0; i < nq; ++i) {                                            /* :120-146 / :281-300 */
        const int w = qw[i];
        if (w < 0 || w >= d->nwords) continue;
        for (int j = 0; j < d->npost[w]; ++j) {
            OrcKf* k = &d->kf[d->post[w][j]];
            if (loop) {
                if (k->loop_query != qid) {
                    k->loop_words = 0;
                    int connected = 0;
                    for (int c = 0; c < nconn; ++c) if (conn[c] == d->post[w][j]) { connected = 1; break; }
                    if (!connected) { k->loop_query = qid; sharing[ns++] = d->post[w][j]; }
                }
                k->loop_words++;
            } else {
                if (k->reloc_query != qid) { k->reloc_words = 0; k->reloc_query = qid; sharing[ns++] = d->post[w][j]; }
                k->reloc_words++;
            }
        }
    }
    if (!ns) { free(sharing); return 0; }
    int maxCommon = 0;                                                        /* :151-159 */
    for (int i = 0; i < ns; ++i) { const OrcKf* k = &d->kf[sharing[i]]; const int c = loop ? k->loop_words : k->reloc_words; if (c > maxCommon) maxCommon = c; }
    const int minCommon = (int)(maxCommon * 0.8f);                            /* int * float -> float -> int, :161 */
    ScoreKf* sm = (ScoreKf*)malloc(sizeof(ScoreKf) * (size_t)ns);
    int nsm = 0;
    for (int i = 0; i < ns; ++i) {                                            /* :166-182 / :315-327 */
        OrcKf* k = &d->kf[sharing[i]];
        if ((loop ? k->loop_words : k->reloc_words) > minCommon) {
            const float si = (float)l1_score(qw, qv, nq, k->word, k->val, k->nbow);
            if (loop) { k->loop_score = si; if (si >= minScore) { sm[nsm].score = si; sm[nsm++].id = sharing[i]; } }
            else { k->reloc_score = si; sm[nsm].score = si; sm[nsm++].id = sharing[i]; }
        }
    }
    free(sharing);
    if (!nsm) { free(sm); return 0; }
    ScoreKf* acc = (ScoreKf*)malloc(sizeof(ScoreKf) * (size_t)nsm);
    float bestAcc = loop ? minScore : 0.f;                                    /* :189 / :334 */
    for (int i = 0; i < nsm; ++i) {                                           /* :193-224 / :337-364 */
        const OrcKf* k = &d->kf[sm[i].id];
        float best = sm[i].score, a = sm[i].score;
        int bestId = sm[i].id;
        for (int j = 0; j < k->nneigh; ++j) {
            const int nid = k->neigh[j];
            if (nid < 0 || nid >= d->nkf || !d->kf[nid].exists) continue;
            const OrcKf* k2 = &d->kf[nid];
            if (loop) {
                if (k2->loop_query == qid && k2->loop_words > minCommon) {
                    a += k2->loop_score;
                    if (k2->loop_score > best) { bestId = nid; best = k2->loop_score; }
                }
            } else {
                if (k2->reloc_query != qid) continue;
                a += k2->reloc_score;                                         /* also of a neighbour not scored in THIS query, :351 */
                if (k2->reloc_score > best) { bestId = nid; best = k2->reloc_score; }
            }
        }
        acc[i].score = a; acc[i].id = bestId;
        if (a > bestAcc) bestAcc = a;
    }
    const float retain = 0.75f * bestAcc;                                     /* :227 / :368 */
    int n = 0;
    for (int i = 0; i < nsm; ++i)                                             /* :236-251 / :373-392: first occurrence only */
        if (acc[i].score > retain) {
            int dup = 0;
            for (int j = 0; j < i && !dup; ++j) dup = acc[j].score > retain && acc[j].id == acc[i].id;
            if (!dup) { if (n < cap) out[n] = acc[i].id; ++n; }
        }
    free(sm); free(acc);
    return n;
}

int orc_kfdb_detect_loop(orc_kfdb* d, int query_kf, const int32_t* conn, int nconn, float minScore, int32_t* out, int cap)
{
    OrcKf* q = kf_slot(d, query_kf);
    if (!q || !q->exists) return -1;
    return detect(d, 1, query_kf, q->word, q->val, q->nbow, conn, nconn, minScore, out, cap);
}

int orc_kfdb_detect_reloc(orc_kfdb* d, long long frame_id, const int32_t* word, const double* val, int n, int32_t* out, int cap)
{
    return detect(d, 0, frame_id, word, val, n, NULL, 0, 0.f, out, cap);
}
