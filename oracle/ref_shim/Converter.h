// Stand-in for the reference's include/Converter.h (which needs Eigen and g2o, absent here): declares only the member
// the compiled translation units (src/Frame.cc:424, src/KeyFrame.cc:63) call.  Defined in oracle/ref_driver_slam.cc.
// TEST INFRASTRUCTURE ONLY.
#ifndef CONVERTER_H
#define CONVERTER_H
#include <opencv2/core/core.hpp>
#include <vector>
namespace ORB_SLAM2 {
class Converter {
 public:
  // one 1 x 32 row header per descriptor, sharing the matrix' buffer
  static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};
}  // namespace ORB_SLAM2
#endif
