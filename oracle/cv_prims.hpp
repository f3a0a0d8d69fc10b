// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
//
// CPU restatement of the OpenCV primitives the reference's ORB front-end calls.
// OpenCV is an un-vendored, un-pinned system dependency of the reference
// (`find_package(OpenCV REQUIRED)`, /root/reference/CMakeLists.txt:13); its source
// is not under /root/reference.  The arithmetic below is the published behaviour of
// OpenCV 4.13.0 (the only OpenCV on this image, as the `cv2` wheel) and every
// function here is pinned bit-for-bit against that wheel by
// tests/test_oracle_vs_cv2.py.
//
// Reference call sites that these stand in for:
//   cv::resize(INTER_LINEAR)      Features/ORBextractor.cpp:1676
//   cv::copyMakeBorder(REFLECT101) Features/ORBextractor.cpp:1689,1708
//   cv::FAST(roi, kps, th, true)  Features/ORBextractor.cpp:1109,1119
//   cv::fastAtan2                 Features/ORBextractor.cpp:158
//   cvRound / cvFloor / cvCeil    Features/ORBextractor.cpp:109,177,187,542,578-586,1663
//   cv::GaussianBlur(7x7, s=2)    Features/ORBextractor.cpp:1601
//   cv::undistortPoints           SlamTypes/Frame.cpp:119,150
//   cv::cvtColor(BGR2GRAY)        demo/demo_initialization.cpp:67-68 (row N1, "next")
//
// Build with -ffp-contract=off and no -march flags: x86-64 baseline has no FMA, and
// the float formulas here rely on separate multiply/add roundings.
#pragma once
#include <cfloat>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <vector>

namespace orc {

typedef unsigned char u8;

// ---- rounding helpers (OpenCV fast_math.hpp semantics on x86-64 SSE2) -------------
#if defined(__SSE2__)
static inline int cv_round(double v) { return _mm_cvtsd_si32(_mm_set_sd(v)); }  // as OpenCV: cvtsd2si, half-to-even
static inline int cv_round(float v) { return _mm_cvtss_si32(_mm_set_ss(v)); }   // cvtss2si
#else
static inline int cv_round(double v) { return (int)std::lrint(v); }
static inline int cv_round(float v) { return (int)std::lrintf(v); }
#endif
static inline int cv_floor(float v) { int i = (int)v; return i - (i > v); }
static inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(float v) { int i = (int)v; return i + (i < v); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (i < v); }

// ---- BORDER_REFLECT_101 index (gfedcb|abcdefgh|gfedcba) ----------------------------
static inline int reflect101(int i, int len) {
  if (len == 1) return 0;
  while (i < 0 || i >= len) {
    if (i < 0) i = -i;
    else i = 2 * (len - 1) - i;
  }
  return i;
}

// ---- cv::copyMakeBorder(src, dst, b,b,b,b, BORDER_REFLECT_101) ------------------------
// dst is (w+2b) x (h+2b).
static inline void copy_make_border101(const u8* src, int w, int h, size_t sstep, u8* dst,
                                       size_t dstep, int b) {
  for (int y = -b; y < h + b; ++y) {
    const u8* srow = src + (size_t)reflect101(y, h) * sstep;
    u8* drow = dst + (size_t)(y + b) * dstep;
    for (int x = -b; x < w + b; ++x) drow[x + b] = srow[reflect101(x, w)];
  }
}

// ---- cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR), CV_8UC1 -----------------------
// Fixed-point bilinear: 11-bit horizontal/vertical coefficients (INTER_RESIZE_COEF_BITS),
// int32 row accumulators, the VResizeLinear<uchar> shift sequence.
struct ResizeTab {
  std::vector<int> ofs;      // source index per destination index
  std::vector<short> coef;   // 2 coefficients per destination index
};
static inline ResizeTab resize_tab_x(int ssize, int dsize) {
  ResizeTab t;
  t.ofs.resize(dsize);
  t.coef.resize(2 * (size_t)dsize);
  double inv_scale = (double)dsize / ssize;
  double scale = 1. / inv_scale;
  for (int d = 0; d < dsize; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = cv_floor(f);
    f -= s;
    if (s < 0) { f = 0; s = 0; }
    if (s >= ssize - 1) { f = 0; s = ssize - 1; }
    t.ofs[d] = s;
    t.coef[2 * d] = (short)cv_round((1.f - f) * 2048.f);
    t.coef[2 * d + 1] = (short)cv_round(f * 2048.f);
  }
  return t;
}
static inline ResizeTab resize_tab_y(int ssize, int dsize) {
  // rows: the fractional part is kept, the two source rows are clipped instead
  ResizeTab t;
  t.ofs.resize(dsize);
  t.coef.resize(2 * (size_t)dsize);
  double inv_scale = (double)dsize / ssize;
  double scale = 1. / inv_scale;
  for (int d = 0; d < dsize; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = cv_floor(f);
    f -= s;
    t.ofs[d] = s;
    t.coef[2 * d] = (short)cv_round((1.f - f) * 2048.f);
    t.coef[2 * d + 1] = (short)cv_round(f * 2048.f);
  }
  return t;
}
static inline int clipi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static inline void resize_linear_u8(const u8* src, int sw, int sh, size_t sstep, u8* dst, int dw,
                                    int dh, size_t dstep) {
  ResizeTab tx = resize_tab_x(sw, dw), ty = resize_tab_y(sh, dh);
  std::vector<int> r0(dw), r1(dw);
  for (int dy = 0; dy < dh; ++dy) {
    int sy0 = clipi(ty.ofs[dy], 0, sh - 1), sy1 = clipi(ty.ofs[dy] + 1, 0, sh - 1);
    const u8* s0 = src + (size_t)sy0 * sstep;
    const u8* s1 = src + (size_t)sy1 * sstep;
    for (int dx = 0; dx < dw; ++dx) {
      int sx = tx.ofs[dx];
      int sx1 = sx + 1 < sw ? sx + 1 : sw - 1;
      int a0 = tx.coef[2 * dx], a1 = tx.coef[2 * dx + 1];
      r0[dx] = s0[sx] * a0 + s0[sx1] * a1;
      r1[dx] = s1[sx] * a0 + s1[sx1] * a1;
    }
    int b0 = ty.coef[2 * dy], b1 = ty.coef[2 * dy + 1];
    u8* d = dst + (size_t)dy * dstep;
    for (int dx = 0; dx < dw; ++dx)
      d[dx] = (u8)((((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2);
  }
}

// ---- cv::FAST(img, kps, threshold, nonmaxSuppression=true), TYPE_9_16 ------------------
struct FastPt { int x, y, score; };

static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// Segment-test value m(p) = max(A, B); A = max over the 16 arcs of 9 of min(v - ring),
// B = same for (ring - v).  p is a corner at threshold t iff m > t; its NMS score is m - 1
// (cornerScore<16> returns max(t, A, B) - 1, which is m - 1 for every corner).
// Returns 0 early when the opposing-pair test proves m <= t (the same rejection cv::FAST does).
static inline int fast_arc_value(const u8* p, const ptrdiff_t* ofs, int t) {
  int d[32];
  const int v = p[0];
  // a 9-arc contains one pixel of every opposing pair: if for some pair both |d| <= t, m <= t
  static const int order[8] = {0, 4, 2, 6, 1, 3, 5, 7};
  for (int i = 0; i < 8; ++i) {
    const int k = order[i];
    const int a = v - p[ofs[k]], b = v - p[ofs[k + 8]];
    if (a <= t && a >= -t && b <= t && b >= -t) return 0;
    d[k] = a; d[k + 8] = b;
  }
  for (int k = 16; k < 32; ++k) d[k] = d[k - 16];
  int mn2[24], mx2[24], mn4[20], mx4[20];
  for (int k = 0; k < 24; ++k) { mn2[k] = d[k] < d[k + 1] ? d[k] : d[k + 1]; mx2[k] = d[k] > d[k + 1] ? d[k] : d[k + 1]; }
  for (int k = 0; k < 20; ++k) { mn4[k] = mn2[k] < mn2[k + 2] ? mn2[k] : mn2[k + 2]; mx4[k] = mx2[k] > mx2[k + 2] ? mx2[k] : mx2[k + 2]; }
  int A = -256, B = -256;
  for (int k = 0; k < 16; ++k) {
    int mn = mn4[k] < mn4[k + 4] ? mn4[k] : mn4[k + 4];
    int mx = mx4[k] > mx4[k + 4] ? mx4[k] : mx4[k + 4];
    if (d[k + 8] < mn) mn = d[k + 8];
    if (d[k + 8] > mx) mx = d[k + 8];
    if (mn > A) A = mn;
    if (-mx > B) B = -mx;
  }
  return A > B ? A : B;
}

// img: cols x rows ROI with row stride `step`.  Output in row-major (y, then x) order,
// coordinates relative to the ROI, exactly as cv::FAST returns them.
static inline void fast9_16_nms(const u8* img, int cols, int rows, size_t step, int threshold,
                                std::vector<FastPt>& out) {
  out.clear();
  if (threshold < 0) threshold = 0;
  if (threshold > 255) threshold = 255;
  if (cols < 7 || rows < 7) return;
  // score rows: 0 where not a corner / outside the tested interior
  std::vector<int> buf((size_t)3 * cols, 0);
  int* rowbuf[3] = {&buf[0], &buf[(size_t)cols], &buf[(size_t)2 * cols]};
  std::vector<std::vector<int>> cpos(3);
  ptrdiff_t ofs[16];
  for (int k = 0; k < 16; ++k) ofs[k] = (ptrdiff_t)kRingDy[k] * (ptrdiff_t)step + kRingDx[k];
  for (int i = 3; i < rows - 2; ++i) {
    int* curr = rowbuf[(i - 3) % 3];
    std::vector<int>& ccur = cpos[(i - 3) % 3];
    std::memset(curr, 0, sizeof(int) * (size_t)cols);
    ccur.clear();
    if (i < rows - 3) {
      const u8* row = img + (size_t)i * step;
      for (int j = 3; j < cols - 3; ++j) {
        int m = fast_arc_value(row + j, ofs, threshold);
        if (m > threshold) {
          ccur.push_back(j);
          curr[j] = m - 1;
        }
      }
    }
    if (i == 3) continue;
    const int* prev = rowbuf[(i - 4 + 3) % 3];
    const int* pprev = rowbuf[(i - 5 + 3) % 3];
    const std::vector<int>& cprev = cpos[(i - 4 + 3) % 3];
    for (size_t k = 0; k < cprev.size(); ++k) {
      int j = cprev[k];
      int s = prev[j];
      if (s > prev[j + 1] && s > prev[j - 1] && s > pprev[j - 1] && s > pprev[j] &&
          s > pprev[j + 1] && s > curr[j - 1] && s > curr[j] && s > curr[j + 1]) {
        FastPt p = {j, i - 1, s};
        out.push_back(p);
      }
    }
  }
}

// ---- cv::fastAtan2(y, x) (degrees) -------------------------------------------------
static inline float fast_atan2(float y, float x) {
  const float scale = (float)(180 / 3.1415926535897932384626433832795);
  const float p1 = 0.9997878412794807f * scale;
  const float p3 = -0.3258083974640975f * scale;
  const float p5 = 0.1555786518463281f * scale;
  const float p7 = -0.04432655554792128f * scale;
  float ax = std::fabs(x), ay = std::fabs(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// ---- cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101), CV_8UC1 -------------
// Fixed-point separable kernel {18,34,48,56,48,34,18}/256 (8.8 horizontal, 8.16 total).
static inline void gaussian_blur7_u8(const u8* src, int w, int h, size_t sstep, u8* dst,
                                     size_t dstep) {
  // row pass into a ring of 7 u16 rows (reflected source rows), column pass out of it
  std::vector<u8> prow((size_t)w + 6);
  std::vector<unsigned short> ring((size_t)7 * w);
  auto hrow = [&](int y, unsigned short* out) {
    const u8* s = src + (size_t)reflect101(y, h) * sstep;
    for (int x = -3; x < w + 3; ++x) prow[(size_t)(x + 3)] = s[reflect101(x, w)];
    const u8* p = prow.data();
    for (int x = 0; x < w; ++x)
      out[x] = (unsigned short)(18 * (p[x] + p[x + 6]) + 34 * (p[x + 1] + p[x + 5]) + 48 * (p[x + 2] + p[x + 4]) +
                                56 * p[x + 3]);
  };
  for (int y = -3; y < 3; ++y) hrow(y, &ring[(size_t)((y + 7) % 7) * w]);
  for (int y = 0; y < h; ++y) {
    hrow(y + 3, &ring[(size_t)((y + 3) % 7) * w]);
    const unsigned short* r[7];
    for (int i = 0; i < 7; ++i) r[i] = &ring[(size_t)((y + i - 3 + 7) % 7) * w];
    u8* d = dst + (size_t)y * dstep;
    for (int x = 0; x < w; ++x) {
      unsigned acc = 18u * (r[0][x] + r[6][x]) + 34u * (r[1][x] + r[5][x]) + 48u * (r[2][x] + r[4][x]) + 56u * r[3][x];
      d[x] = (u8)((acc + 32768u) >> 16);
    }
  }
}

// ---- cv::undistortPoints(pts, pts, K, D=[k1,k2,p1,p2], noArray(), P=K) ---------------
// K = (fx, fy, cx, cy) and D are float32 matrices widened to double; 5 fixed-point
// iterations of the inverse Brown model; result re-projected with K and narrowed to float.
static inline void undistort_point(float u, float v, const float K[4], const float D[4],
                                   float* ou, float* ov) {
  double fx = K[0], fy = K[1], cx = K[2], cy = K[3];
  double k1 = D[0], k2 = D[1], p1 = D[2], p2 = D[3];
  double ifx = 1. / fx, ify = 1. / fy;
  double x = ((double)u - cx) * ifx, y = ((double)v - cy) * ify;
  double x0 = x, y0 = y;
  for (int j = 0; j < 5; ++j) {
    double r2 = x * x + y * y;
    double icdist = 1. / (1 + ((0 * r2 + k2) * r2 + k1) * r2);
    if (icdist < 0) {  // OpenCV 4.x (regression test 14583): a negative radial factor -- far outside the model's range -- gives up
      x = ((double)u - cx) * ifx;
      y = ((double)v - cy) * ify;
      break;
    }
    double dx = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
    double dy = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
    x = (x0 - dx) * icdist;
    y = (y0 - dy) * icdist;
  }
  *ou = (float)(x * fx + cx);
  *ov = (float)(y * fy + cy);
}

// ---- cv::cvtColor(BGR2GRAY), CV_8UC3 -> CV_8UC1 ---------------------------------------
static inline u8 bgr2gray(u8 b, u8 g, u8 r) {
  return (u8)((3735 * b + 19235 * g + 9798 * r + 16384) >> 15);
}

// cv::KeyPointsFilter::retainBest(keypoints, n_points) -- OpenCV (un-vendored dependency of the reference; semantics of
// the 4.x line, modules/features2d/src/keypoint.cpp, restated; the Python wheel does not expose it, so unlike the other
// primitives it cannot be pinned against cv2 here: PARITY UNPINNED for this one function): when more than n_points are
// held, std::nth_element puts the n_points largest responses first, every later keypoint whose response ties the n-th one
// is kept as well (std::partition), the rest is dropped.  The element order is whatever the C++ library's nth_element and
// partition leave -- libstdc++ here, as for std::sort in DistributeOctTree.  KP needs a float member `response`.
template <typename KP>
static inline void retain_best(std::vector<KP>& k, int n_points) {
  if (n_points >= 0 && k.size() > (size_t)n_points) {
    if (n_points == 0) { k.clear(); return; }
    std::nth_element(k.begin(), k.begin() + n_points - 1, k.end(), [](const KP& a, const KP& b) { return a.response > b.response; });
    const float ambiguous = k[n_points - 1].response;
    typename std::vector<KP>::iterator e = std::partition(k.begin() + n_points, k.end(), [ambiguous](const KP& a) { return a.response >= ambiguous; });
    k.resize((size_t)(e - k.begin()));
  }
}

}  // namespace orc
