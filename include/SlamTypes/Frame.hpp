// Drop-in for the reference's SlamTypes/Frame.hpp (front-end part: constructor glue, grid,
// GetFeaturesInArea; reference SlamTypes/Frame.hpp:20-136, Frame.cpp:40-206).  Members that belong to
// out-of-scope subsystems (BoW vectors, map points, pose) are omitted here; a full integration keeps
// the reference's own header and replaces only the bodies shown in INTEGRATION.md.
#pragma once
#include <cmath>
#include <vector>

#include "../Features/ORBextractor.hpp"

namespace ORB_SLAM_Tracking {

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class Frame {
 public:
  Frame() : mpORBextractor(nullptr), mTimestamp(0), N(0), mnId(0) {}
  // K: 3x3 CV_32F, distCoef: 4x1 CV_32F (k1,k2,p1,p2) as in the reference.  voc is unused (nullptr in the demo).
  Frame(cv::Mat& im_, const double& timestamp, ORBextractor* extractor, void* /*voc*/, cv::Mat& K, cv::Mat& distCoef)
      : mpORBextractor(extractor), im(im_), mTimestamp(timestamp), mK(K), mDistCoef(distCoef), N(0) {
    orbx_ctx* ctx = extractor->context(im.cols, im.rows);
    const float K4[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    const float D4[4] = {distCoef.at<float>(0), distCoef.at<float>(1), distCoef.at<float>(2), distCoef.at<float>(3)};
    if (mbInitialComputations()) {  // once per process, Frame.cpp:44-55
      int b[4];
      orbx_set_camera(ctx, K4, D4, nullptr, b);
      mnMinX() = b[0]; mnMaxX() = b[1]; mnMinY() = b[2]; mnMaxY() = b[3];
      mfGridElementWidthInv() = (float)FRAME_GRID_COLS / (float)(b[1] - b[0]);
      mfGridElementHeightInv() = (float)FRAME_GRID_ROWS / (float)(b[3] - b[2]);
      fx() = K4[0]; fy() = K4[1]; cx() = K4[2]; cy() = K4[3];
      mbInitialComputations() = false;
    } else {
      const int b[4] = {mnMinX(), mnMaxX(), mnMinY(), mnMaxY()};
      orbx_set_camera(ctx, K4, D4, b, nullptr);
    }
    std::vector<int> lap = {0, 0};
    (*mpORBextractor)(im, cv::Mat(), mvKeys, mDescriptors, lap);  // Frame.cpp:58-59
    N = (int)mvKeys.size();
    if (mvKeys.empty()) return;  // :60-61
    mvKeysUn.resize(N);
    gridOff_.resize(ORBX_GRID_CELLS + 1);
    gridIdx_.resize(N);
    orbx_frame_glue(ctx, reinterpret_cast<const orbx_keypoint*>(mvKeys.data()), N,
                    reinterpret_cast<orbx_keypoint*>(mvKeysUn.data()), gridOff_.data(), gridIdx_.data());
    mnId = nNextId()++;
    for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
      for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
        const int c = ix * FRAME_GRID_ROWS + iy;
        mGrid[ix][iy].assign(gridIdx_.begin() + gridOff_[c], gridIdx_.begin() + gridOff_[c + 1]);
      }
  }

  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY) {  // Frame.cpp:89-99
    posX = (int)std::round((kp.pt.x - mnMinX()) * mfGridElementWidthInv());
    posY = (int)std::round((kp.pt.y - mnMinY()) * mfGridElementHeightInv());
    return !(posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS);
  }

  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                        const int maxLevel = -1) const {
    std::vector<size_t> out;
    if (!mpORBextractor || N == 0) return out;
    std::vector<int> tmp(N);
    const orbx_frame f = pod();
    const int n = orbx_features_in_area(mpORBextractor->context(), &f, x, y, r, minLevel, maxLevel, tmp.data(), N);
    for (int i = 0; i < n && i < N; ++i) out.push_back((size_t)tmp[i]);
    return out;
  }

  orbx_frame pod() const {
    orbx_frame f;
    f.n = N;
    f.keys_un = reinterpret_cast<const orbx_keypoint*>(mvKeysUn.data());
    f.desc = mDescriptors.data;
    f.grid_offsets = gridOff_.data();
    f.grid_indices = gridIdx_.data();
    return f;
  }

  ORBextractor* mpORBextractor;
  cv::Mat im;
  double mTimestamp;
  cv::Mat mK, mDistCoef;
  int N;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  cv::Mat mDescriptors;
  long unsigned int mnId;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

  // statics of the reference (Frame.cpp:4-8), header-only here
  static float& fx() { static float v; return v; }
  static float& fy() { static float v; return v; }
  static float& cx() { static float v; return v; }
  static float& cy() { static float v; return v; }
  static long unsigned int& nNextId() { static long unsigned int v = 0; return v; }
  static int& mnMinX() { static int v; return v; }
  static int& mnMaxX() { static int v; return v; }
  static int& mnMinY() { static int v; return v; }
  static int& mnMaxY() { static int v; return v; }
  static float& mfGridElementWidthInv() { static float v; return v; }
  static float& mfGridElementHeightInv() { static float v; return v; }
  static bool& mbInitialComputations() { static bool v = true; return v; }

 private:
  std::vector<uint32_t> gridOff_, gridIdx_;
};

}  // namespace ORB_SLAM_Tracking
