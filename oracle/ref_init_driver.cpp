// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
//
// Pins the oracle's restatement of Initializer::CheckHomography / CheckFundamental against the reference's own
// function bodies.  Initialization/Initializer.cpp as a whole cannot be compiled here (cv::findHomography,
// decomposeEssentialMat, triangulatePoints, cv::Mat algebra, Eigen: none installed), so oracle/build_ref.sh cuts
// the constructor and the two scoring functions out of the file where it lies under /root/reference into
// oracle/_ref/initializer_slice.inc (generated at build time, git-ignored, never committed) and this file compiles
// them unmodified against the reference's own Initializer.hpp.
#include <cstring>
#include <vector>

#define private public  // the scoring functions and the members they read are private (Initializer.hpp:57-199)
#include "Initialization/Initializer.hpp"
#undef private

namespace ORB_SLAM_Tracking {
#include "initializer_slice.inc"
}  // namespace ORB_SLAM_Tracking

using namespace ORB_SLAM_Tracking;

namespace {
void fill(Initializer& I, const cv::KeyPoint* k2, int n2, const int* m12, int n1) {
  I.mvKeys2.assign(k2, k2 + n2);
  I.mvMatches12.clear();
  for (int i = 0; i < n1; ++i)  // Initializer::Initialize, Initializer.cpp:27-34
    if (m12[i] >= 0) I.mvMatches12.push_back(std::make_pair(i, m12[i]));
}
Eigen::Matrix3f mat(const float* p) { Eigen::Matrix3f M; std::memcpy(M.m, p, sizeof(M.m)); return M; }
}  // namespace

extern "C" {

int ref_check_homography(const void* k1, int n1, const void* k2, int n2, const int* m12, const float* H21s, const float* H12s,
                         int nHyp, float sigma, float* scores, unsigned char* inliers) {
  Frame F;
  F.mK = cv::Mat(3, 3, CV_32F);
  F.mvKeysUn.assign((const cv::KeyPoint*)k1, (const cv::KeyPoint*)k1 + n1);
  Initializer I(F, sigma, nHyp);
  fill(I, (const cv::KeyPoint*)k2, n2, m12, n1);
  const int N = (int)I.mvMatches12.size();
  for (int h = 0; h < nHyp; ++h) {
    std::vector<bool> in;
    scores[h] = I.CheckHomography(mat(H21s + 9 * h), mat(H12s + 9 * h), in, sigma);
    if (inliers) for (int i = 0; i < N; ++i) inliers[(size_t)h * N + i] = in[i] ? 1 : 0;
  }
  return N;
}

int ref_check_fundamental(const void* k1, int n1, const void* k2, int n2, const int* m12, const float* F21s, int nHyp, float sigma,
                          float* scores, unsigned char* inliers) {
  Frame F;
  F.mK = cv::Mat(3, 3, CV_32F);
  F.mvKeysUn.assign((const cv::KeyPoint*)k1, (const cv::KeyPoint*)k1 + n1);
  Initializer I(F, sigma, nHyp);
  fill(I, (const cv::KeyPoint*)k2, n2, m12, n1);
  const int N = (int)I.mvMatches12.size();
  for (int h = 0; h < nHyp; ++h) {
    std::vector<bool> in;
    scores[h] = I.CheckFundamental(mat(F21s + 9 * h), in, sigma);
    if (inliers) for (int i = 0; i < N; ++i) inliers[(size_t)h * N + i] = in[i] ? 1 : 0;
  }
  return N;
}

}  // extern "C"
