// Drop-in for the reference's Features/ORBextractor.hpp (ORB_SLAM_Tracking::ORBextractor,
// reference Features/ORBextractor.hpp:55-158): same class name, namespace, constructor,
// operator() (5 parameters, returns monoIndex), accessors and the public mvImagePyramid member.
// Host C++ only; every computation goes through the C ABI in orbx.h to the CUDA library.
// The reference binds an extractor to no image size; this one creates its device context lazily
// for the first image's size (and again if the size changes).
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <list>
#include <vector>

#include "../cv_compat.hpp"
#include "../orbx.h"

namespace ORB_SLAM_Tracking {

namespace orbx_detail {
#ifdef ORBX_HAVE_OPENCV
inline cv::Mat toMat(cv::InputArray a) { return a.getMat(); }
inline void createOut(cv::OutputArray o, int r, int c, int t) { o.create(r, c, t); }
inline cv::Mat outMat(cv::OutputArray o) { return o.getMat(); }
inline void releaseOut(cv::OutputArray o) { o.release(); }
#else
inline cv::Mat toMat(cv::InputArray a) { return a; }
inline void createOut(cv::OutputArray o, int r, int c, int t) { o.create(r, c, t); }
inline cv::Mat outMat(cv::OutputArray o) { return o; }
inline void releaseOut(cv::OutputArray o) { o.release(); }
#endif
}  // namespace orbx_detail

// Public type of the reference (Features/ORBextractor.hpp:32-53), kept for source compatibility with the same members.
// The extractor's own quadtree runs on the device (csrc/quadtree.cu) and does not use it.
class ExtractorNode {
 public:
  ExtractorNode() : bNoMore(false) {}

  // Features/ORBextractor.cpp:617-676: four children at ceil(w / 2), ceil(h / 2); a key goes left / top when its
  // coordinate is below the split line; a child with exactly one key is final.
  void DivideNode(ExtractorNode& n1, ExtractorNode& n2, ExtractorNode& n3, ExtractorNode& n4) {
    const int hx = (int)std::ceil(static_cast<float>(UR.x - UL.x) / 2), hy = (int)std::ceil(static_cast<float>(BR.y - UL.y) / 2);
    const cv::Point2i mid(UL.x + hx, UL.y + hy);
    n1.UL = UL;                         n1.UR = cv::Point2i(mid.x, UL.y); n1.BL = cv::Point2i(UL.x, mid.y);  n1.BR = mid;
    n2.UL = n1.UR;                      n2.UR = UR;                       n2.BL = mid;                       n2.BR = cv::Point2i(UR.x, mid.y);
    n3.UL = n1.BL;                      n3.UR = mid;                      n3.BL = BL;                        n3.BR = cv::Point2i(mid.x, BL.y);
    n4.UL = mid;                        n4.UR = n2.BR;                    n4.BL = n3.BR;                     n4.BR = BR;
    ExtractorNode* child[4] = {&n1, &n2, &n3, &n4};
    for (ExtractorNode* c : child) c->vKeys.reserve(vKeys.size());
    for (const cv::KeyPoint& kp : vKeys) child[(kp.pt.x < mid.x ? 0 : 1) + (kp.pt.y < mid.y ? 0 : 2)]->vKeys.push_back(kp);
    for (ExtractorNode* c : child)
      if (c->vKeys.size() == 1) c->bNoMore = true;
  }

  std::vector<cv::KeyPoint> vKeys;
  cv::Point2i UL, UR, BL, BR;
  std::list<ExtractorNode>::iterator lit;
  bool bNoMore;
};

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_, int device = 0,
               int maxBatch = 1)
      : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_),
        minThFAST(minThFAST_), device_(device), maxBatch_(maxBatch), ctx_(nullptr), w_(0), h_(0) {
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
    int umax[16];
    const int rc = orbx_compute_level_tables(nfeatures, scaleFactor_, nlevels, mvScaleFactor.data(),
                                             mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                                             mnFeaturesPerLevel.data(), umax);
    if (rc == ORBX_ERR_CTOR) {  // reference: message + exit(1), Features/ORBextractor.cpp:502-505
      std::fprintf(stderr, "scaleFactor == 1.0 but nlevels is not 1\n");
      std::exit(1);
    }
    int sum = 0;
    for (int v : mnFeaturesPerLevel) sum += v;
    std::cout << "Sum of features = " << sum << std::endl;  // the reference's line, :549
  }
  ~ORBextractor() { orbx_destroy(ctx_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Features/ORBextractor.cpp:1531-1653.  `mask` is ignored, as in the reference (:1537).
  int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                 cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
    cv::Mat image = orbx_detail::toMat(_image);
    if (image.empty()) return -1;  // :1536
    if (image.type() != CV_8UC1) {  // the reference asserts, :1541
      std::fprintf(stderr, "ORBextractor::operator(): assertion `image.type() == CV_8UC1' failed (Features/ORBextractor.cpp:1541)\n");
      std::abort();
    }
    if (!ensure(image.cols, image.rows)) return -1;
    const int cap = capacity();
    _keypoints.resize(cap);
    std::vector<uint8_t> desc((size_t)cap * 32);
    const int lap[2] = {vLappingArea.size() > 0 ? vLappingArea[0] : 0, vLappingArea.size() > 1 ? vLappingArea[1] : 0};
    int n = 0, mono = 0;
    const int rc = orbx_extract(ctx_, image.data, image.step, lap, reinterpret_cast<orbx_keypoint*>(_keypoints.data()),
                                desc.data(), cap, &n, &mono);
    if (rc != ORBX_OK) { std::fprintf(stderr, "orbx_extract: %d %s\n", rc, orbx_last_error(ctx_)); _keypoints.clear(); return -1; }
    _keypoints.resize(n);
    pyramidStale_ = true;
    if (fillPyramid_) FetchImagePyramid();
    if (n == 0) { orbx_detail::releaseOut(_descriptors); return 0; }  // :1567-1570 (mono index of an empty result)
    orbx_detail::createOut(_descriptors, n, 32, CV_8U);
    cv::Mat d = orbx_detail::outMat(_descriptors);
    for (int i = 0; i < n; ++i) std::memcpy(d.ptr<uint8_t>(i), &desc[(size_t)i * 32], 32);
    return mono;
  }
  // north_star's 4-argument form: no lapping area
  int operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors) {
    std::vector<int> lap = {0, 0};
    return (*this)(image, mask, keypoints, descriptors, lap);
  }

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }
  std::vector<int> inline GetNumFeaturesPerLevel() { return mnFeaturesPerLevel; }

  // Public member of the reference (hpp:111): the levels of the last image (unpadded views, as the reference's ROIs).
  // Filled after every operator() like in the reference; a caller that never reads it can switch the device -> host
  // copy off (SetFillImagePyramid(false)) and fetch on demand with FetchImagePyramid().
  std::vector<cv::Mat> mvImagePyramid;
  void SetFillImagePyramid(bool on) { fillPyramid_ = on; }
  const std::vector<cv::Mat>& FetchImagePyramid() {
    if (ctx_ && pyramidStale_) {
      mvImagePyramid.resize(nlevels);
      for (int l = 0; l < nlevels; ++l) {
        int w = 0, h = 0;
        orbx_get_level_size(ctx_, l, &w, &h);
        mvImagePyramid[l].create(h, w, CV_8U);
        orbx_get_pyramid_level(ctx_, 0, l, 0, mvImagePyramid[l].data, mvImagePyramid[l].step);
      }
      pyramidStale_ = false;
    }
    return mvImagePyramid;
  }

  // Frame::Frame extracts through orbx_frame_construct on this extractor's context: same bookkeeping as operator()
  void NotifyExtracted() {
    pyramidStale_ = true;
    if (fillPyramid_) FetchImagePyramid();
  }
  orbx_ctx* context(int width, int height) { return ensure(width, height) ? ctx_ : nullptr; }
  orbx_ctx* context() { return ctx_; }
  int capacity() const {
    int s = 0;
    for (int v : mnFeaturesPerLevel) s += v;
    return s > nfeatures ? s : nfeatures;
  }

 protected:
  bool ensure(int w, int h) {
    if (ctx_ && w == w_ && h == h_) return true;
    orbx_destroy(ctx_);
    int st = 0;
    ctx_ = orbx_create(device_, w, h, nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, maxBatch_, &st);
    if (!ctx_) { std::fprintf(stderr, "orbx_create(%dx%d) failed: %d\n", w, h, st); return false; }
    w_ = w; h_ = h;
    return true;
  }

  int nfeatures;
  double scaleFactor;  // double holding the float, as in the reference (hpp:142)
  int nlevels, iniThFAST, minThFAST;
  std::vector<int> mnFeaturesPerLevel;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  int device_, maxBatch_;
  orbx_ctx* ctx_;
  int w_, h_;
  bool pyramidStale_ = false, fillPyramid_ = true;
};

}  // namespace ORB_SLAM_Tracking
