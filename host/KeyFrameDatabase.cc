// Host side of the KeyFrameDatabase drop-in (include/KeyFrameDatabase.h): flattens the SLAM objects for orbx_kfdb_*
// (csrc/k_kfdb.cu) and turns candidate ids back into KeyFrame pointers.  Reference: src/KeyFrameDatabase.cc.
#include "KeyFrameDatabase.h"
#include "orbx.h"
#include <cstdio>
#include <cstdlib>

namespace ORB_SLAM2
{

static void flatten(const DBoW2::BowVector& b, std::vector<int32_t>& w, std::vector<double>& v)
{
    w.clear(); v.clear();
    w.reserve(b.size()); v.reserve(b.size());
    for (DBoW2::BowVector::const_iterator it = b.begin(); it != b.end(); ++it) { w.push_back((int32_t)it->first); v.push_back(it->second); }
}

void KeyFrameDatabase::Init(int nwords) { mnWords = nwords; }

void KeyFrameDatabase::Note(int rc, const char* what)
{
    mnLastStatus = rc;
    if (rc && !mbReported) { fprintf(stderr, "KeyFrameDatabase (orbx): %s failed: %s\n", what, orbx_strerror(rc)); mbReported = true; }
    if (!rc) mbReported = false;
}

bool KeyFrameDatabase::Ready()
{
    if (mpDb) return true;
    int dev = 0;
    if (const char* e = getenv("ORBX_DEVICE")) dev = atoi(e);
    int rc = mpCtx ? ORBX_OK : orbx_create_matcher(&mpCtx, dev);
    if (!rc) rc = orbx_kfdb_create(mpCtx, mnWords > 0 ? mnWords : 1, &mpDb);
    if (rc) { Note(rc, "orbx_kfdb_create"); return false; }
    return true;
}

KeyFrameDatabase::~KeyFrameDatabase()
{
    if (mpDb) orbx_kfdb_destroy(mpDb);
    if (mpCtx) orbx_destroy(mpCtx);
}

void KeyFrameDatabase::add(KeyFrame* pKF)
{
    std::unique_lock<std::mutex> lock(mMutex);
    if (!pKF || !Ready()) return;
    std::vector<int32_t> w; std::vector<double> v;
    flatten(pKF->mBowVec, w, v);
    const int rc = orbx_kfdb_add(mpDb, (int)pKF->mnId, w.data(), v.data(), (int)w.size());
    Note(rc, "orbx_kfdb_add");
    if (!rc) mKeyFrames[pKF->mnId] = pKF;
}

void KeyFrameDatabase::erase(KeyFrame* pKF)
{
    std::unique_lock<std::mutex> lock(mMutex);
    if (!pKF || !Ready()) return;
    Note(orbx_kfdb_erase(mpDb, (int)pKF->mnId), "orbx_kfdb_erase");
    mKeyFrames.erase(pKF->mnId);
}

void KeyFrameDatabase::clear()
{
    std::unique_lock<std::mutex> lock(mMutex);
    if (!Ready()) return;
    Note(orbx_kfdb_clear(mpDb), "orbx_kfdb_clear");
    mKeyFrames.clear();
}

void KeyFrameDatabase::SetCovisible(KeyFrame* pKF)
{
    std::unique_lock<std::mutex> lock(mMutex);
    if (!pKF || !Ready() || !mKeyFrames.count(pKF->mnId)) return;
    const std::vector<KeyFrame*> best = pKF->GetBestCovisibilityKeyFrames(10);
    std::vector<int32_t> ids;
    for (size_t j = 0; j < best.size(); ++j) ids.push_back((int32_t)best[j]->mnId);
    Note(orbx_kfdb_set_covisible(mpDb, (int)pKF->mnId, ids.data(), (int)ids.size()), "orbx_kfdb_set_covisible");
}

void KeyFrameDatabase::RefreshCovisibility()
{
    if (!mbRefresh) return;
    const int n = (int)mKeyFrames.size();
    std::vector<int32_t> ids((size_t)n), nb((size_t)n * 10, -1);
    int i = 0;
    for (std::map<long unsigned int, KeyFrame*>::iterator it = mKeyFrames.begin(); it != mKeyFrames.end(); ++it, ++i) {
        ids[i] = (int32_t)it->first;
        const std::vector<KeyFrame*> best = it->second->GetBestCovisibilityKeyFrames(10);
        for (size_t j = 0; j < best.size() && j < 10; ++j) nb[(size_t)i * 10 + j] = (int32_t)best[j]->mnId;
    }
    Note(orbx_kfdb_set_covisible_bulk(mpDb, n, ids.data(), nb.data()), "orbx_kfdb_set_covisible_bulk");
}

std::vector<KeyFrame*> KeyFrameDatabase::DetectLoopCandidates(KeyFrame* pKF, float minScore)
{
    std::vector<KeyFrame*> out;
    if (!pKF) return out;
    const std::set<KeyFrame*> sp = pKF->GetConnectedKeyFrames();      // taken outside the database mutex, like :116
    std::unique_lock<std::mutex> lock(mMutex);
    if (!Ready() || mKeyFrames.empty()) return out;
    RefreshCovisibility();
    if (mnLastStatus) return out;
    std::vector<int32_t> w, conn; std::vector<double> v;
    flatten(pKF->mBowVec, w, v);
    for (std::set<KeyFrame*>::const_iterator it = sp.begin(); it != sp.end(); ++it) conn.push_back((int32_t)(*it)->mnId);
    std::vector<int32_t> ids(mKeyFrames.size());
    int n = 0;
    const int rc = orbx_kfdb_detect_loop(mpDb, (long long)pKF->mnId, w.data(), v.data(), (int)w.size(), conn.data(), (int)conn.size(), minScore,
                                         ids.data(), (int)ids.size(), &n);
    Note(rc, "orbx_kfdb_detect_loop");
    if (rc) return out;
    out.reserve((size_t)n);
    for (int i = 0; i < n && i < (int)ids.size(); ++i) {
        // a candidate is a held keyframe or one of its covisible neighbours that a query met -- both are in the map
        std::map<long unsigned int, KeyFrame*>::iterator it = mKeyFrames.find((long unsigned int)ids[i]);
        if (it != mKeyFrames.end()) out.push_back(it->second);
    }
    return out;
}

std::vector<KeyFrame*> KeyFrameDatabase::DetectRelocalizationCandidates(Frame* F)
{
    std::vector<KeyFrame*> out;
    if (!F) return out;
    std::unique_lock<std::mutex> lock(mMutex);
    if (!Ready() || mKeyFrames.empty()) return out;
    RefreshCovisibility();
    if (mnLastStatus) return out;
    std::vector<int32_t> w; std::vector<double> v;
    flatten(F->mBowVec, w, v);
    const long long qid = (long long)F->mnId;
    const int32_t seg[2] = {0, (int32_t)w.size()};
    std::vector<int32_t> ids(mKeyFrames.size());
    int n = 0;
    const int rc = orbx_kfdb_detect_reloc(mpDb, 1, &qid, w.data(), v.data(), seg, ids.data(), (int)ids.size(), &n);
    Note(rc, "orbx_kfdb_detect_reloc");
    if (rc) return out;
    out.reserve((size_t)n);
    for (int i = 0; i < n && i < (int)ids.size(); ++i) {
        std::map<long unsigned int, KeyFrame*>::iterator it = mKeyFrames.find((long unsigned int)ids[i]);
        if (it != mKeyFrames.end()) out.push_back(it->second);
    }
    return out;
}

} // namespace ORB_SLAM2
