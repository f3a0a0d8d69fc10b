// Minimal stand-ins for the OpenCV types the reference's front-end API exposes, used ONLY when
// <opencv2/core.hpp> is not on the include path (this build box has no OpenCV C++ headers).
// With OpenCV present the real headers are used and this file is empty.  Layouts match OpenCV:
// cv::KeyPoint is 28 bytes (pt.x, pt.y, size, angle, response, octave, class_id).
#pragma once
#if defined(__has_include)
#if __has_include(<opencv2/core.hpp>)
#define ORBX_HAVE_OPENCV 1
#endif
#endif

#ifdef ORBX_HAVE_OPENCV
#include <opencv2/core.hpp>
#else
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

namespace cv {

struct Point2f {
  float x, y;
  Point2f() : x(0), y(0) {}
  Point2f(float x_, float y_) : x(x_), y(y_) {}
};

struct Point2i {
  int x, y;
  Point2i() : x(0), y(0) {}
  Point2i(int x_, int y_) : x(x_), y(y_) {}
};
typedef Point2i Point;

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

// 2-D, single channel, reference-counted; enough for images (CV_8U), descriptor matrices
// (CV_8U n x 32) and the 3x3 / 4x1 CV_32F camera matrices.
class Mat {
 public:
  int rows, cols;
  size_t step;
  unsigned char* data;
  Mat() : rows(0), cols(0), step(0), data(nullptr), type_(CV_8U) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(int r, int c, int type, void* ext, size_t step_ = 0) : rows(r), cols(c), data((unsigned char*)ext), type_(type) {
    step = step_ ? step_ : (size_t)c * elemSize();
  }
  void create(int r, int c, int type) {
    if (r == rows && c == cols && type == type_ && data && own_) return;
    rows = r; cols = c; type_ = type; step = (size_t)c * elemSize();
    own_.reset(new unsigned char[(size_t)r * step + 1], std::default_delete<unsigned char[]>());
    data = own_.get();
  }
  void release() { own_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  size_t elemSize() const { return type_ == CV_32F ? 4 : 1; }
  bool isContinuous() const { return step == (size_t)cols * elemSize(); }
  template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
  template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
  template <typename T> T& at(int r, int c = 0) { return ptr<T>(r)[c]; }
  template <typename T> const T& at(int r, int c = 0) const { return ptr<T>(r)[c]; }
  Mat row(int r) const { Mat m = *this; m.rows = 1; m.data = data + (size_t)r * step; return m; }
  Mat clone() const {
    Mat m(rows, cols, type_);
    for (int r = 0; r < rows; ++r) std::memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols * elemSize());
    return m;
  }

 private:
  int type_;
  std::shared_ptr<unsigned char> own_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#endif
