// Test driver for the C++ drop-in class: reads a raw 8-bit image, runs ORB_SLAM2::ORBextractor::operator() exactly
// the way Frame::ExtractORB does (src/Frame.cc:430-443), and dumps keypoints, descriptors and mvImagePyramid.
//   orbx_cpp_driver <in.raw> <w> <h> <nfeatures> <out.bin>
#include "ORBextractor.h"
#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char** argv)
{
    if (argc < 6) { fprintf(stderr, "usage\n"); return 2; }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]);
    std::vector<unsigned char> buf((size_t)w * h);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) { fprintf(stderr, "read failed\n"); return 2; }
    fclose(f);
    try {
        ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
        cv::Mat im(h, w, CV_8UC1, buf.data());
        std::vector<cv::KeyPoint> keys;
        cv::Mat desc;
        ex(im, cv::Mat(), keys, desc);
        ex(im, cv::Mat(), keys, desc); // second call: context reuse
        if (ex.LastStatus() != 0) {       // the class itself never throws; the driver reports what it recorded
            fprintf(stderr, "ORBextractor::LastStatus() = %d, %d keypoints, descriptors %s\n", ex.LastStatus(), (int)keys.size(), desc.empty() ? "released" : "kept");
            return 3;
        }
        FILE* o = fopen(argv[5], "wb");
        int n = (int)keys.size(), L = ex.GetLevels();
        fwrite(&n, 4, 1, o);
        fwrite(keys.data(), sizeof(cv::KeyPoint), n, o);
        for (int i = 0; i < n; ++i) fwrite(desc.ptr<unsigned char>(i), 1, 32, o);
        fwrite(&L, 4, 1, o);
        std::vector<float> sf = ex.GetScaleFactors();
        fwrite(sf.data(), 4, L, o);
        for (int l = 0; l < L; ++l) {
            const cv::Mat& m = ex.mvImagePyramid[l];
            fwrite(&m.cols, 4, 1, o); fwrite(&m.rows, 4, 1, o);
            // the ROI sits inside its bordered parent: dump with the 19-px frame, like Frame.cc reads around keypoints
            for (int y = -19; y < m.rows + 19; ++y) fwrite(m.data + (long)y * (long)m.step - 19, 1, m.cols + 38, o);
        }
        fclose(o);
    } catch (const std::exception& e) { fprintf(stderr, "error: %s\n", e.what()); return 1; }
    return 0;
}
