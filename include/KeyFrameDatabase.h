// Drop-in replacement for the reference's include/KeyFrameDatabase.h (QiuYue-bit/ORB-SLAM2-noted, :44-87): same namespace,
// class and method signatures, so LoopClosing (src/LoopClosing.cc: DetectLoop -> DetectLoopCandidates), Tracking
// (src/Tracking.cc: Relocalization -> DetectRelocalizationCandidates), KeyFrame::SetBadFlag (erase) and System::Reset (clear)
// call it unchanged.  The inverted-file walk and the scoring run on the GPU behind the C ABI (include/orbx.h: orbx_kfdb_*);
// host/KeyFrameDatabase.cc only flattens BowVectors and covisibility lists and maps ids back to KeyFrame pointers.
#ifndef KEYFRAMEDATABASE_H
#define KEYFRAMEDATABASE_H

#include <list>
#include <map>
#include <mutex>
#include <set>
#include <vector>

#include "KeyFrame.h"
#include "Frame.h"

struct orbx_ctx;
struct orbx_kfdb;

namespace ORB_SLAM2
{

class KeyFrame;
class Frame;

class KeyFrameDatabase
{
public:
    // The reference takes `const ORBVocabulary& voc` and reads voc.size() (:38-42); any vocabulary type with size() works.
    template <class Vocabulary> explicit KeyFrameDatabase(const Vocabulary& voc) { Init((int)voc.size()); }
    ~KeyFrameDatabase();

    void add(KeyFrame* pKF);                                                        // src/KeyFrameDatabase.cc:50-56
    void erase(KeyFrame* pKF);                                                      // :64-83
    void clear();                                                                   // :85-89
    std::vector<KeyFrame*> DetectLoopCandidates(KeyFrame* pKF, float minScore);     // :113-270
    std::vector<KeyFrame*> DetectRelocalizationCandidates(Frame* F);                // :275-399

    // ---- additions (not in the reference) ----
    // Like the reference, nothing throws.  A library error (no device, CUDA failure) makes a query return no candidates,
    // prints one line to stderr and is kept here (0 = ORBX_OK) until a call succeeds.
    int LastStatus() const { return mnLastStatus; }
    // The queries need every held keyframe's GetBestCovisibilityKeyFrames(10).  By default they are all re-read in front of a
    // query (one upload).  A tree that calls SetCovisible(this) from KeyFrame::UpdateBestCovisibles can turn that off.
    void SetCovisible(KeyFrame* pKF);
    void SetCovisibilityRefresh(bool on) { mbRefresh = on; }
    // Number of words of a size()-less vocabulary stand-in
    struct Words { int n; size_t size() const { return (size_t)n; } };

protected:
    void Init(int nwords);
    bool Ready();
    void Note(int rc, const char* what);
    void RefreshCovisibility();     // every held keyframe's GetBestCovisibilityKeyFrames(10), in one upload, in front of a query
    int mnWords = 0;
    orbx_ctx* mpCtx = nullptr;
    orbx_kfdb* mpDb = nullptr;
    std::map<long unsigned int, KeyFrame*> mKeyFrames;      // id -> keyframe, for the way back from candidate ids
    std::mutex mMutex;                                      // the reference's mMutex (:84): one caller at a time
    int mnLastStatus = 0;
    bool mbReported = false;
    bool mbRefresh = true;
};

} // namespace ORB_SLAM2

#endif
