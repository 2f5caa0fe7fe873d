// ref_driver.h — shared between ref_driver.cc and ref_driver_slam.cc (TEST INFRASTRUCTURE ONLY).
#ifndef ORACLE_REF_DRIVER_H_
#define ORACLE_REF_DRIVER_H_
#include <vector>

#include "ORBextractor.h"

struct ref_extractor {
  ORB_SLAM2::ORBextractor* ex;
  std::vector<cv::KeyPoint> kps;  // results of the last ref_extract
  cv::Mat desc;
};

// one reference call: mutes the reference's std::cout lines and (deterministic mode) turns the monotonic list-node
// arena on for every thread the call spawns
struct RefCallScope {
  RefCallScope();
  ~RefCallScope();
  void* mute_;
};
#endif
