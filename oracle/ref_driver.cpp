// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
//
// C entry points around the reference's own classes, linked with the reference's own sources
// (compiled where they lie under /root/reference by oracle/build_ref.sh) into
// oracle/_ref/libref_orb.so.  Used by tests/test_oracle_vs_ref.py and tools/gen_golden.py to pin
// the restatement in oracle/orb_oracle.cpp; never by the product path.
#include <cstdio>
#include <cstring>
#include <atomic>
#include <thread>
#include <vector>

#include "Features/ORBextractor.hpp"
#include "Features/ORBmatcher.hpp"
#include "SlamTypes/Frame.hpp"

using namespace ORB_SLAM_Tracking;

namespace {
struct Quiet {  // the reference prints counters to std::cout (ORBextractor.cpp:549, ORBmatcher.cpp:144-147)
  std::streambuf* old;
  Quiet() : old(std::cout.rdbuf(nullptr)) {}
  ~Quiet() { std::cout.rdbuf(old); }
};
}  // namespace

extern "C" {

void* ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  Quiet q;
  return new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_extractor_destroy(void* e) { delete (ORBextractor*)e; }

void ref_extractor_tables(void* e_, float* scale, float* invScale, float* sigma2, float* invSigma2, int* perLevel) {
  ORBextractor* e = (ORBextractor*)e_;
  const int n = e->GetLevels();
  std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
                     d = e->GetInverseScaleSigmaSquares();
  std::vector<int> f = e->GetNumFeaturesPerLevel();
  for (int i = 0; i < n; ++i) { scale[i] = a[i]; invScale[i] = b[i]; sigma2[i] = c[i]; invSigma2[i] = d[i]; perLevel[i] = f[i]; }
}

// ORBextractor::operator() on a host image; returns the keypoint count (or -1), writes monoIndex
int ref_extract(void* e_, const unsigned char* img, int w, int h, int step, int lap0, int lap1, void* kps, unsigned char* desc,
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
  static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
  if (n) std::memcpy(kps, keys.data(), (size_t)n * 28);
  for (int i = 0; i < n; ++i) std::memcpy(desc + (size_t)i * 32, d.ptr<unsigned char>(i), 32);
  return n;
}

// padded = 0: the level itself; 1: with the 19-px border (the parent buffer of mvImagePyramid[level])
int ref_get_level(void* e_, int level, int padded, unsigned char* dst, int* w, int* h) {
  ORBextractor* e = (ORBextractor*)e_;
  if (level < 0 || level >= (int)e->mvImagePyramid.size()) return -1;
  const cv::Mat& m = e->mvImagePyramid[level];
  const int b = padded ? 19 : 0;
  *w = m.cols + 2 * b;
  *h = m.rows + 2 * b;
  if (dst)
    for (int y = -b; y < m.rows + b; ++y) std::memcpy(dst + (size_t)(y + b) * *w, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - b, (size_t)*w);
  return 0;
}

// Frame::Frame(im, ts, extractor, nullptr, K, distCoef).  first != 0 re-arms the once-per-process
// static initialisation (Frame.cpp:44-55), as a new process with this camera would see it.
void* ref_frame_create(void* e_, const unsigned char* img, int w, int h, int step, const float K4[4], const float D4[4], int first) {
  cv::Mat im(h, w, CV_8UC1, (void*)img, (size_t)step);
  cv::Mat K = cv::Mat::zeros(3, 3, CV_32F);
  K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3]; K.at<float>(2, 2) = 1.f;
  cv::Mat D(4, 1, CV_32F);
  for (int i = 0; i < 4; ++i) D.at<float>(i) = D4[i];
  if (first) Frame::mbInitialComputations = true;
  return new Frame(im, 0.0, (ORBextractor*)e_, nullptr, K, D);
}
void ref_frame_destroy(void* f) { delete (Frame*)f; }
int ref_frame_n(void* f) { return ((Frame*)f)->N; }
void ref_frame_bounds(int b[4]) { b[0] = Frame::mnMinX; b[1] = Frame::mnMaxX; b[2] = Frame::mnMinY; b[3] = Frame::mnMaxY; }
void ref_frame_get(void* f_, void* keys, void* keysUn, unsigned char* desc) {
  Frame* f = (Frame*)f_;
  if (f->mvKeys.empty()) return;
  std::memcpy(keys, f->mvKeys.data(), f->mvKeys.size() * 28);
  std::memcpy(keysUn, f->mvKeysUn.data(), f->mvKeysUn.size() * 28);
  for (int i = 0; i < f->N; ++i) std::memcpy(desc + (size_t)i * 32, f->mDescriptors.ptr<unsigned char>(i), 32);
}
// CSR copy of mGrid in cell order x * 48 + y (the order GetFeaturesInArea walks it)
void ref_frame_grid(void* f_, unsigned* off, unsigned* idx) {
  Frame* f = (Frame*)f_;
  unsigned n = 0;
  for (int x = 0; x < FRAME_GRID_COLS; ++x)
    for (int y = 0; y < FRAME_GRID_ROWS; ++y) {
      off[x * FRAME_GRID_ROWS + y] = n;
      for (size_t i : f->mGrid[x][y]) idx[n++] = (unsigned)i;
    }
  off[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
}
int ref_features_in_area(void* f_, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap) {
  std::vector<size_t> v = ((Frame*)f_)->GetFeaturesInArea(x, y, r, minLevel, maxLevel);
  for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int)v[i];
  return (int)v.size();
}
int ref_search_for_initialization(void* f1, void* f2, float nnratio, int checkOri, int window, int* matches12) {
  Quiet q;
  ORBmatcher m(nnratio, checkOri != 0);
  std::vector<int> v;
  const int n = m.SearchForInitialization(*(Frame*)f1, *(Frame*)f2, v, window);
  for (size_t i = 0; i < v.size(); ++i) matches12[i] = v[i];
  return n;
}
double ref_forb_distance(const unsigned char* a, const unsigned char* b) {
  cv::Mat A(1, 32, CV_8U, (void*)a), B(1, 32, CV_8U, (void*)b);
  return DBoW2::FORB::distance(A, B);
}

// CPU baseline: the reference's own Frame construction + SearchForInitialization over a stream of
// frames, pairs (2k, 2k+1) sharded over `threads` host threads (one ORBextractor per thread: the
// reference's extractor is not re-entrant).  Returns the total number of matches; *totalKp = keypoints.
long long ref_run_stream(const unsigned char* imgs, int n, int w, int h, int nfeatures, float scaleFactor, int nlevels, int iniTh,
                         int minTh, const float K4[4], const float D4[4], int window, float nnratio, int threads,
                         long long* totalKp) {
  Quiet q;
  cv::Mat K = cv::Mat::zeros(3, 3, CV_32F);
  K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3]; K.at<float>(2, 2) = 1.f;
  cv::Mat D(4, 1, CV_32F);
  for (int i = 0; i < 4; ++i) D.at<float>(i) = D4[i];
  {  // the once-per-process statics (Frame.cpp:44-55), before any thread starts
    ORBextractor e(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    cv::Mat im(h, w, CV_8UC1, (void*)imgs, (size_t)w);
    Frame::mbInitialComputations = true;
    Frame f(im, 0.0, &e, nullptr, K, D);
  }
  const int nPairs = n / 2;
  std::atomic<int> next(0);
  std::atomic<long long> matches(0), kps(0);
  auto work = [&]() {
    ORBextractor e(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    ORBmatcher m(nnratio, true);
    for (;;) {
      const int p = next.fetch_add(1);
      if (p >= nPairs) break;
      cv::Mat a(h, w, CV_8UC1, (void*)(imgs + (size_t)(2 * p) * w * h), (size_t)w);
      cv::Mat b(h, w, CV_8UC1, (void*)(imgs + (size_t)(2 * p + 1) * w * h), (size_t)w);
      cv::Mat Ka = K, Da = D;
      Frame f1(a, 0.0, &e, nullptr, Ka, Da), f2(b, 1.0, &e, nullptr, Ka, Da);
      std::vector<int> v;
      matches += m.SearchForInitialization(f1, f2, v, window);
      kps += f1.N + f2.N;
    }
  };
  std::vector<std::thread> T;
  for (int t = 1; t < threads; ++t) T.emplace_back(work);
  work();
  for (auto& t : T) t.join();
  if (totalKp) *totalKp = kps.load();
  return matches.load();
}
}
