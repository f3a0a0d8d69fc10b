// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
//
// A stand-in for the OpenCV C++ headers, just large enough to compile the reference's OWN hot-path
// sources (/root/reference/Features/ORBextractor.cpp, Features/ORBmatcher.cpp, SlamTypes/Frame.cpp,
// Thirdparty/DBoW2/src/FORB.cpp) from where they lie, into oracle/_ref/libref_orb.so.  OpenCV C++
// is not installed on this image; the OpenCV *primitives* the reference calls (resize, copyMakeBorder,
// FAST, GaussianBlur, fastAtan2, undistortPoints, cvRound...) are served by oracle/cv_prims.hpp, which
// tests/test_oracle_vs_cv2.py pins bit-for-bit against the real OpenCV 4.13.0 wheel.  Everything
// else that runs in libref_orb.so -- the cell loop, DivideNode/DistributeOctTree with the real
// std::list and std::sort, IC_Angle, computeOrbDescriptor, output packing, Frame's grid,
// GetFeaturesInArea, SearchForInitialization, ComputeThreeMaxima, FORB::distance -- is the
// reference's unmodified code.  tests/test_oracle_vs_ref.py checks the restatement in
// oracle/orb_oracle.cpp against it.
#pragma once
#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "../../cv_prims.hpp"

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_PI 3.1415926535897932384626433832795
#define CV_CN_SHIFT 3
#define CV_8U 0
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_Assert(expr) assert(expr)

static inline int cvRound(double v) { return orc::cv_round(v); }
static inline int cvRound(float v) { return orc::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return orc::cv_floor(v); }
static inline int cvFloor(float v) { return orc::cv_floor(v); }
static inline int cvCeil(double v) { return orc::cv_ceil(v); }
static inline int cvCeil(float v) { return orc::cv_ceil(v); }

namespace cv {

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
  template <typename U>
  Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
  Point_& operator*=(float s) {  // OpenCV: saturate_cast<T>(a.x * b) with float b
    x = (T)(x * s);
    y = (T)(y * s);
    return *this;
  }
  Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
template <typename T>
struct Point3_ {
  T x, y, z;
  Point3_() : x(0), y(0), z(0) {}
  Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;  // Initialization/Initializer.hpp signatures

template <typename T>
struct Size_ {
  T width, height;
  Size_() : width(0), height(0) {}
  Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T>
struct Rect_ {
  T x, y, width, height;
  Rect_() : x(0), y(0), width(0), height(0) {}
  Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Scalar {
  double val[4];
  Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; }
};

struct Range {
  int start, end;
  Range(int s, int e) : start(s), end(e) {}
};

class KeyPoint {  // 28 bytes, the layout of cv::KeyPoint
 public:
  Point2f pt;
  float size;
  float angle;
  float response;
  int octave;
  int class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(Point2f p, float s, float a = -1, float r = 0, int o = 0, int c = -1)
      : pt(p), size(s), angle(a), response(r), octave(o), class_id(c) {}
  KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
      : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};

enum BorderTypes { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3,
                   BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1 };

static inline size_t elem_size_of(int type) {
  static const size_t d[7] = {1, 1, 2, 2, 4, 4, 8};
  return d[type & 7] * (size_t)((type >> CV_CN_SHIFT) + 1);
}

class Mat {
 public:
  int rows, cols;
  uchar* data;
  size_t step;  // bytes per row

  Mat() : rows(0), cols(0), data(nullptr), step(0), type_(0) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, void* ext, size_t st = 0)  // borrowed memory
      : rows(r), cols(c), data((uchar*)ext), step(st ? st : (size_t)c * elem_size_of(type)), type_(type) {}

  void create(int r, int c, int type) {
    if (data && r == rows && c == cols && type == type_) return;  // as cv::Mat::create: no-op when it fits
    type_ = type;
    rows = r;
    cols = c;
    step = (size_t)c * elem_size_of(type);
    buf_ = std::shared_ptr<uchar>(new uchar[std::max<size_t>(step * (size_t)r, 1)], std::default_delete<uchar[]>());
    data = buf_.get();
  }
  void create(Size sz, int type) { create(sz.height, sz.width, type); }
  void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  int depth() const { return type_ & 7; }
  int channels() const { return (type_ >> CV_CN_SHIFT) + 1; }
  size_t elemSize() const { return elem_size_of(type_); }
  Size size() const { return Size(cols, rows); }
  size_t total() const { return (size_t)rows * cols; }
  size_t step1() const { return step / elem_size_of(type_ & 7); }
  bool isContinuous() const { return step == (size_t)cols * elemSize() || rows <= 1; }

  static Mat zeros(int r, int c, int type) {
    Mat m(r, c, type);
    for (int y = 0; y < r; ++y) std::memset(m.data + (size_t)y * m.step, 0, (size_t)c * m.elemSize());
    return m;
  }
  Mat clone() const {
    Mat m;
    copyTo(m);
    return m;
  }
  void copyTo(Mat& dst) const {
    dst.create(rows, cols, type_);
    for (int y = 0; y < rows; ++y) std::memmove(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * elemSize());
  }
  void copyTo(Mat&& dst) const { copyTo(dst); }  // desc.row(i).copyTo(descriptors.row(j)): the row header fits, so it writes in place
  void convertTo(Mat& dst, int rtype) const {
    assert(depth() == CV_8U && (rtype & 7) == CV_32F);
    Mat out(rows, cols, CV_MAKETYPE(CV_32F, channels()));
    for (int y = 0; y < rows; ++y)
      for (int x = 0; x < cols * channels(); ++x) ((float*)(out.data + (size_t)y * out.step))[x] = (float)(data + (size_t)y * step)[x];
    dst = out;
  }
  Mat reshape(int cn) const {  // same rows, channel count changed
    Mat m = *this;
    const int d = depth();
    m.cols = cols * channels() / cn;
    m.type_ = CV_MAKETYPE(d, cn);
    return m;
  }

  Mat operator()(const Rect& r) const {
    assert(r.x >= 0 && r.y >= 0 && r.x + r.width <= cols && r.y + r.height <= rows);
    Mat m = *this;
    m.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
    m.rows = r.height;
    m.cols = r.width;
    return m;
  }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return rowRange(y, y + 1); }

  template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step))[x]; }
  template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step))[x]; }
  template <typename T> T& at(int i) { return rows == 1 ? ((T*)data)[i] : *(T*)(data + (size_t)i * step); }
  template <typename T> const T& at(int i) const { return rows == 1 ? ((const T*)data)[i] : *(const T*)(data + (size_t)i * step); }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  uchar* ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }

 private:
  int type_;
  std::shared_ptr<uchar> buf_;
};

class _InputArray {
 public:
  _InputArray() : m_(nullptr) {}
  _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
  bool empty() const { return !m_ || m_->empty(); }
  Mat getMat() const { return m_ ? *m_ : Mat(); }
 protected:
  Mat* m_;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray(Mat& m) : _InputArray(m) {}
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  void create(Size s, int type) const { m_->create(s, type); }
  void release() const { m_->release(); }
  Mat& getMatRef() const { return *m_; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
static inline InputArray noArray() { static _InputArray none; return none; }

// ---- the primitives, served by oracle/cv_prims.hpp (pinned against the cv2 4.13.0 wheel) ------------

static inline float fastAtan2(float y, float x) { return orc::fast_atan2(y, x); }

static inline void resize(InputArray src_, OutputArray dst_, Size dsize, double fx = 0, double fy = 0,
                          int interpolation = INTER_LINEAR) {
  (void)fx; (void)fy;
  assert(interpolation == INTER_LINEAR);
  Mat src = src_.getMat();
  assert(src.type() == CV_8UC1 && dsize.width > 0 && dsize.height > 0);
  dst_.create(dsize, src.type());
  Mat dst = dst_.getMat();
  orc::resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

static inline void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right,
                                  int borderType, const Scalar& = Scalar()) {
  // The reference only pads whole images or (with BORDER_ISOLATED) the centre ROI of dst itself.
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && left == right && top == left);
  Mat src = src_.getMat().clone();  // src may alias the centre of dst
  assert(src.type() == CV_8UC1);
  dst_.create(src.rows + top + bottom, src.cols + left + right, src.type());
  Mat dst = dst_.getMat();
  orc::copy_make_border101(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top);
}

static inline void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY = 0,
                                int borderType = BORDER_DEFAULT) {
  assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
  Mat src = src_.getMat().clone();  // in-place call in the reference
  assert(src.type() == CV_8UC1);
  dst_.create(src.rows, src.cols, src.type());
  Mat dst = dst_.getMat();
  orc::gaussian_blur7_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

static inline void FAST(InputArray image_, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true) {
  assert(nonmaxSuppression);
  Mat img = image_.getMat();
  assert(img.type() == CV_8UC1);
  std::vector<orc::FastPt> pts;
  orc::fast9_16_nms(img.data, img.cols, img.rows, img.step, threshold, pts);
  keypoints.clear();
  for (const orc::FastPt& p : pts) keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1.f, (float)p.score));
}

static inline void undistortPoints(InputArray src_, OutputArray dst_, InputArray K_, InputArray D_, InputArray R_ = noArray(),
                                   InputArray P_ = noArray()) {
  Mat src = src_.getMat(), K = K_.getMat(), D = D_.getMat(), P = P_.getMat();
  assert(R_.empty() && src.type() == CV_MAKETYPE(CV_32F, 2) && K.type() == CV_32FC1 && D.type() == CV_32FC1);
  assert(P.data == K.data);  // the reference passes P = K
  assert(D.total() == 4);
  const float K4[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
  const float D4[4] = {D.at<float>(0), D.at<float>(1), D.at<float>(2), D.at<float>(3)};
  dst_.create(src.rows, src.cols, src.type());
  Mat dst = dst_.getMat();
  for (int y = 0; y < src.rows; ++y)
    for (int x = 0; x < src.cols; ++x) {
      const float* s = src.ptr<float>(y) + 2 * x;
      float* d = dst.ptr<float>(y) + 2 * x;
      float u, v;
      orc::undistort_point(s[0], s[1], K4, D4, &u, &v);
      d[0] = u;
      d[1] = v;
    }
}

struct KeyPointsFilter {
  // only called from the reference's dead ComputeKeyPointsOld path (Features/ORBextractor.cpp:1452,1482)
  static void retainBest(std::vector<KeyPoint>& k, int n) { orc::retain_best(k, n); }
};

}  // namespace cv
