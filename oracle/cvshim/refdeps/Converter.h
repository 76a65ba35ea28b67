// Stand-in for the reference's include/Converter.h, which needs Eigen and g2o headers (external, absent here).
// Only what src/Frame.cc and src/KeyFrame.cc call: Converter::toDescriptorVector (include/Converter.h:38,
// src/Converter.cc:29-38 — one 1 x 32 row header per descriptor).  ORACLE support, test infrastructure only.
#ifndef CONVERTER_H
#define CONVERTER_H
#include <vector>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors) {
        std::vector<cv::Mat> v;
        v.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) v.push_back(Descriptors.row(j));
        return v;
    }
};
}
#endif
