// Drop-in for the reference's Features/ORBextractor.hpp (ORB_SLAM_Tracking::ORBextractor,
// reference Features/ORBextractor.hpp:55-158): same class name, namespace, constructor,
// operator() (5 parameters, returns monoIndex), accessors and the public mvImagePyramid member.
// Host C++ only; every computation goes through the C ABI in orbx.h to the CUDA library.
// The reference binds an extractor to no image size; this one creates its device context lazily
// for the first image's size (and again if the size changes).
#pragma once
#include <cstdio>
#include <cstdlib>
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

// Declaration kept for source compatibility (reference Features/ORBextractor.hpp:32-53); the
// quadtree itself runs on the device (csrc/quadtree.cu).
class ExtractorNode {
 public:
  ExtractorNode() : bNoMore(false) {}
  std::vector<cv::KeyPoint> vKeys;
  cv::Point2f UL, UR, BL, BR;
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
    std::printf("Sum of features: %d\n", sum);  // :549
  }
  ~ORBextractor() { orbx_destroy(ctx_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Features/ORBextractor.cpp:1531-1653.  `mask` is ignored, as in the reference (:1537).
  int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                 cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
    cv::Mat image = orbx_detail::toMat(_image);
    if (image.empty()) return -1;  // :1536
    if (image.type() != CV_8UC1) { std::fprintf(stderr, "ORBextractor: image must be CV_8UC1\n"); std::abort(); }  // assert, :1541
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

  // Public member of the reference (hpp:111).  Filled on request from the device copy.
  std::vector<cv::Mat> mvImagePyramid;
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
  bool pyramidStale_ = false;
};

}  // namespace ORB_SLAM_Tracking
