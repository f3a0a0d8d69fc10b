// Drop-in for the reference's Features/ORBmatcher.hpp (reference Features/ORBmatcher.hpp:7-63,
// ORBmatcher.cpp:11-150): same class, constructor and SearchForInitialization signature, plus the
// north_star's DescriptorDistance (== DBoW2::FORB::distance, Thirdparty/DBoW2/src/FORB.cpp:77-100).
#pragma once
#include <cstdio>
#include <iostream>
#include <vector>

#include "../SlamTypes/Frame.hpp"

namespace ORB_SLAM_Tracking {

class ORBmatcher {
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

  int SearchForInitialization(Frame& F1, Frame& F2, std::vector<int>& vnMatches12, int windowSize = 100) {
    vnMatches12.assign(F1.N, -1);
    if (!F1.mpORBextractor || !F1.mpORBextractor->context()) return 0;
    const orbx_frame f1 = F1.pod(), f2 = F2.pod();
    int nmatches = 0, cnt[3] = {0, 0, 0};
    const int rc = orbx_search_for_initialization(F1.mpORBextractor->context(), &f1, &f2, mfNNratio, mbCheckOrientation ? 1 : 0,
                                                  windowSize, vnMatches12.data(), &nmatches, cnt);
    if (rc != ORBX_OK) { std::fprintf(stderr, "orbx_search_for_initialization: %d\n", rc); return 0; }
    // the reference's statistics lines, ORBmatcher.cpp:144-147
    std::cout << "SearchForInitialization done ----------------------" << std::endl;
    std::cout << "invalidMatchByDistance: " << cnt[0] << std::endl;
    std::cout << "invalidMatchByRatio: " << cnt[1] << std::endl;
    std::cout << "invalidMatchByOrientation: " << cnt[2] << std::endl;
    return nmatches;
  }

  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbx_descriptor_distance(a.data, b.data); }

  static const int HISTO_LENGTH = 30;
  static const int TH_LOW = 50;
  static const int TH_HIGH = 100;

 private:
  float mfNNratio;
  bool mbCheckOrientation;
};

}  // namespace ORB_SLAM_Tracking
