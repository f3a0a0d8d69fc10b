// Drop-in for the reference's SlamTypes/Frame.hpp (reference SlamTypes/Frame.hpp:20-136, Frame.cpp:40-206): same public
// members, statics as DATA members (Frame::fx ...), copy constructor, SetPose, PosInGrid, GetFeaturesInArea.  The two BoW
// members (mBowVec, mFeatVec) need DBoW2's headers and exist only when ORBX_HAVE_DBOW2 is defined; PoseT is Eigen's
// Affine3d when <Eigen/Geometry> is on the include path (as in SlamTypes/BasicTypes.hpp), else a 4x4 double stand-in.
#pragma once
#include <cmath>
#include <vector>

#include "../Features/ORBextractor.hpp"
#if defined(__has_include)
#if __has_include(<Eigen/Geometry>)
#include <Eigen/Geometry>
#define ORBX_HAVE_EIGEN 1
#endif
#endif
#ifdef ORBX_HAVE_DBOW2
#include "Features/ORBVocabulary.hpp"
#endif

namespace ORB_SLAM_Tracking {

#ifdef ORBX_HAVE_EIGEN
typedef Eigen::Affine3d PoseT;  // SlamTypes/BasicTypes.hpp
#else
struct PoseT { double m[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}; };
#endif
#ifndef ORBX_HAVE_DBOW2
class ORBVocabulary;  // DBoW2::TemplatedVocabulary<...> in the reference; only ever passed around as a pointer (nullptr in the demo)
#endif
class MapPoint;

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class Frame {
 public:
  Frame() : mpORBvocabulary(nullptr), mpORBextractor(nullptr), mTimestamp(0), N(0), mnId(0) {}
  // Frame.cpp:12-38: deep copies of the images / matrices, the pose only if it was set
  Frame(const Frame& f)
      : mpORBvocabulary(f.mpORBvocabulary), mpORBextractor(f.mpORBextractor), im(f.im.clone()), mTimestamp(f.mTimestamp),
        mK(f.mK.clone()), mDistCoef(f.mDistCoef.clone()), N(f.N), mvKeys(f.mvKeys), mvKeysUn(f.mvKeysUn),
        mDescriptors(f.mDescriptors.clone()), mnId(f.mnId), mvpMapPoints(f.mvpMapPoints), mvbOutlier(f.mvbOutlier),
        gridOff_(f.gridOff_), gridIdx_(f.gridIdx_) {
    for (int i = 0; i < FRAME_GRID_COLS; i++)
      for (int j = 0; j < FRAME_GRID_ROWS; j++) mGrid[i][j] = f.mGrid[i][j];
    if (f.mbPoseSet) SetPose(f.mTcw);
  }
  Frame& operator=(const Frame&) = default;
  // K: 3x3 CV_32F, distCoef: 4x1 CV_32F (k1,k2,p1,p2) as in the reference (Frame.cpp:40-82)
  Frame(cv::Mat& im_, const double& timestamp, ORBextractor* extractor, ORBVocabulary* voc, cv::Mat& K, cv::Mat& distCoef)
      : mpORBvocabulary(voc), mpORBextractor(extractor), im(im_), mTimestamp(timestamp), mK(K), mDistCoef(distCoef), N(0) {
    orbx_ctx* ctx = extractor->context(im.cols, im.rows);
    const float K4[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    const float D4[4] = {distCoef.at<float>(0), distCoef.at<float>(1), distCoef.at<float>(2), distCoef.at<float>(3)};
    if (mbInitialComputations) {  // once per process, Frame.cpp:44-55
      int b[4];
      orbx_set_camera(ctx, K4, D4, nullptr, b);
      mnMinX = b[0]; mnMaxX = b[1]; mnMinY = b[2]; mnMaxY = b[3];
      mfGridElementWidthInv = (float)FRAME_GRID_COLS / (float)(b[1] - b[0]);
      mfGridElementHeightInv = (float)FRAME_GRID_ROWS / (float)(b[3] - b[2]);
      fx = K4[0]; fy = K4[1]; cx = K4[2]; cy = K4[3];
      mbInitialComputations = false;
    } else if (ctx != cameraCtx_) {  // a new device context (other image size): hand it the process-wide bounds
      const int b[4] = {mnMinX, mnMaxX, mnMinY, mnMaxY};
      orbx_set_camera(ctx, K4, D4, b, nullptr);
    }
    cameraCtx_ = ctx;
    // Frame.cpp:57-77 in one device round trip: extraction, UndistortKeyPoints and the grid assignment run as one CUDA
    // graph (orbx_frame_construct == (*mpORBextractor)(im, cv::Mat(), mvKeys, mDescriptors, {0, 0}) + the glue)
    if (im.empty() || im.type() != CV_8UC1) {
      std::vector<int> lap = {0, 0};
      (*mpORBextractor)(im, cv::Mat(), mvKeys, mDescriptors, lap);  // the extractor's own error behaviour
      N = 0;
      return;
    }
    const int cap = mpORBextractor->capacity();
    mvKeys.resize(cap); mvKeysUn.resize(cap);
    gridOff_.resize(ORBX_GRID_CELLS + 1); gridIdx_.resize(cap);
    std::vector<uint8_t> desc((size_t)cap * 32);
    int n = 0;
    const int rc = orbx_frame_construct(ctx, im.data, im.step, nullptr, reinterpret_cast<orbx_keypoint*>(mvKeys.data()), desc.data(), cap, &n,
                                        nullptr, reinterpret_cast<orbx_keypoint*>(mvKeysUn.data()), gridOff_.data(), gridIdx_.data());
    if (rc != ORBX_OK) { std::fprintf(stderr, "orbx_frame_construct: %d %s\n", rc, orbx_last_error(ctx)); n = 0; }
    N = n;
    mvKeys.resize(N); mvKeysUn.resize(N); gridIdx_.resize(N);
    mpORBextractor->NotifyExtracted();
    if (N == 0) { mDescriptors.release(); return; }  // :60-61
    mDescriptors.create(N, 32, CV_8U);
    for (int i = 0; i < N; ++i) std::memcpy(mDescriptors.ptr<uint8_t>(i), &desc[(size_t)i * 32], 32);
    mnId = nNextId++;
    for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
      for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
        const int c = ix * FRAME_GRID_ROWS + iy;
        mGrid[ix][iy].assign(gridIdx_.begin() + gridOff_[c], gridIdx_.begin() + gridOff_[c + 1]);
      }
  }

  void SetPose(const PoseT& Tcw) {  // Frame.cpp:84-87
    mTcw = Tcw;
    mbPoseSet = true;
  }

  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY) {  // Frame.cpp:89-99
    posX = (int)std::round((kp.pt.x - mnMinX) * mfGridElementWidthInv);
    posY = (int)std::round((kp.pt.y - mnMinY) * mfGridElementHeightInv);
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

  ORBVocabulary* mpORBvocabulary;
  ORBextractor* mpORBextractor;
  cv::Mat im;
  double mTimestamp;
  cv::Mat mK;
  inline static float fx, fy, cx, cy;  // statics of the reference (Frame.hpp:73, Frame.cpp:4-8), header-only here (C++17)
  cv::Mat mDistCoef;
  int N;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
#ifdef ORBX_HAVE_DBOW2
  DBoW2::BowVector mBowVec;
  DBoW2::FeatureVector mFeatVec;
#endif
  cv::Mat mDescriptors;
  inline static long unsigned int nNextId = 0;
  long unsigned int mnId;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  inline static int mnMinX, mnMaxX, mnMinY, mnMaxY;
  inline static float mfGridElementWidthInv, mfGridElementHeightInv;
  inline static bool mbInitialComputations = true;
  PoseT mTcw;
  bool mbPoseSet = false;

 private:
  std::vector<uint32_t> gridOff_, gridIdx_;   // the grid in CSR form, as the C ABI hands it over (mGrid is built from it)
  inline static orbx_ctx* cameraCtx_ = nullptr;
};

}  // namespace ORB_SLAM_Tracking
