// TEST INFRASTRUCTURE ONLY (oracle/).  Shadows /root/reference/include/Converter.h for the _ref build: the real header
// needs Eigen and g2o, which this image lacks, and the path only uses toDescriptorVector (src/Converter.cc:40-56,
// called at src/Frame.cc:711 and src/KeyFrame.cc:84): one 1x32 row view per descriptor.
#ifndef ORB_REF_CONVERTER_H
#define ORB_REF_CONVERTER_H
#include <opencv2/core/core.hpp>
#include <vector>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors)
    {
        std::vector<cv::Mat> v;
        v.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) v.push_back(Descriptors.row(j));
        return v;
    }
};
}
#endif
