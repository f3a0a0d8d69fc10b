// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
//
// C entry points around the reference's ORBextractor built with its ComputeKeyPointsOld policy (oracle/build_ref.sh
// compiles a generated copy of Features/ORBextractor.cpp whose operator() calls ComputeKeyPointsOld instead of
// ComputeKeyPointsOctTree, under the namespace name ORB_SLAM_Tracking_OldPolicy -- this file is compiled with the same
// -DORB_SLAM_Tracking=ORB_SLAM_Tracking_OldPolicy).  Everything else of operator() (pyramid, blur, descriptors, packing) is
// the reference's unmodified code.
#include <cstring>
#include <iostream>
#include <vector>

#include "Features/ORBextractor.hpp"

using namespace ORB_SLAM_Tracking;  // renamed by the build flag

extern "C" {

void* ref_old_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  std::streambuf* old = std::cout.rdbuf(nullptr);  // "Sum of features = ..."
  void* e = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
  std::cout.rdbuf(old);
  return e;
}
void ref_old_extractor_destroy(void* e) { delete (ORBextractor*)e; }

int ref_old_extract(void* e_, const unsigned char* img, int w, int h, int step, int lap0, int lap1, void* kps, unsigned char* desc,
                    int cap, int* mono) {
  ORBextractor* e = (ORBextractor*)e_;
  cv::Mat im(h, w, CV_8UC1, (void*)img, (size_t)step);
  std::vector<cv::KeyPoint> keys;
  cv::Mat d;
  std::vector<int> lap = {lap0, lap1};
  const int m = (*e)(im, cv::Mat(), keys, d, lap);
  if (mono) *mono = m;
  if (m < 0) return -1;
  const int n = (int)keys.size();
  if (n > cap) return -2;
  if (n) std::memcpy(kps, keys.data(), (size_t)n * 28);
  for (int i = 0; i < n; ++i) std::memcpy(desc + (size_t)i * 32, d.ptr<unsigned char>(i), 32);
  return n;
}

}  // extern "C"
