// Replays a script of KeyFrameDatabase operations through the C++ drop-in (include/KeyFrameDatabase.h) on KeyFrame / Frame
// stand-ins, the way the reference's threads would call it; tests/test_gpu_cpp.py compares the printed candidate lists with
// the oracle's and the reference's.   orbx_cpp_kfdb_driver <script.bin> <nwords>
// Script records (little endian): int32 op, int32 id, int32 n, then
//   0 new  : n x int32 words, n x float64 weights      1 add     2 erase     3 covisible: n x int32 ids (strongest first)
//   4 loop : n x int32 connected ids, float32 minScore  5 reloc: n x int32 words, n x float64 weights     6 clear
#include "KeyFrameDatabase.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>
using namespace ORB_SLAM2;

template <typename T> static bool rd(FILE* f, T* p, size_t n) { return n == 0 || fread(p, sizeof(T), n, f) == n; }

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: orbx_cpp_kfdb_driver script.bin nwords\n"); return 2; }
    FILE* f = fopen(argv[1], "rb");
    if (!f) return 2;
    KeyFrameDatabase::Words voc = {atoi(argv[2])};
    KeyFrameDatabase db(voc);
    std::map<int, KeyFrame*> kfs;
    int32_t hdr[3];
    while (fread(hdr, 4, 3, f) == 3) {
        const int op = hdr[0], id = hdr[1], n = hdr[2];
        std::vector<int32_t> a((size_t)n);
        std::vector<double> v;
        if (op == 0 || op == 3 || op == 4 || op == 5) if (!rd(f, a.data(), (size_t)n)) return 3;
        if (op == 0 || op == 5) { v.resize((size_t)n); if (!rd(f, v.data(), (size_t)n)) return 3; }
        if (op == 0) {
            KeyFrame* k = new KeyFrame();
            k->mnId = (long unsigned)id;
            for (int i = 0; i < n; ++i) k->mBowVec[(unsigned)a[i]] = v[i];
            kfs[id] = k;
        } else if (op == 1) db.add(kfs[id]);
        else if (op == 2) db.erase(kfs[id]);
        else if (op == 6) db.clear();
        else if (op == 3) {
            kfs[id]->mvpOrderedConnectedKeyFrames.clear();
            for (int i = 0; i < n; ++i) if (kfs.count(a[i])) kfs[id]->mvpOrderedConnectedKeyFrames.push_back(kfs[a[i]]);
        } else if (op == 4) {
            float minScore;
            if (!rd(f, &minScore, 1)) return 3;
            kfs[id]->mConnectedKeyFrameWeights.clear();
            for (int i = 0; i < n; ++i) if (kfs.count(a[i])) kfs[id]->mConnectedKeyFrameWeights[kfs[a[i]]] = 15;
            const std::vector<KeyFrame*> c = db.DetectLoopCandidates(kfs[id], minScore);
            printf("loop %d:", id);
            for (size_t i = 0; i < c.size(); ++i) printf(" %lu", c[i]->mnId);
            printf("\n");
        } else if (op == 5) {
            Frame F;
            F.mnId = (long unsigned)id;
            for (int i = 0; i < n; ++i) F.mBowVec[(unsigned)a[i]] = v[i];
            const std::vector<KeyFrame*> c = db.DetectRelocalizationCandidates(&F);
            printf("reloc %d:", id);
            for (size_t i = 0; i < c.size(); ++i) printf(" %lu", c[i]->mnId);
            printf("\n");
        }
    }
    fclose(f);
    if (db.LastStatus()) { fprintf(stderr, "status %d\n", db.LastStatus()); return 4; }
    return 0;
}
