// Exercises the C++ drop-in headers the way demo_initialization.cpp:65-108 uses the reference classes:
// two gray images -> ORBextractor -> two Frames -> ORBmatcher(0.9, true).SearchForInitialization.
// Input: raw u8 file with two W x H images.  Prints N1 N2 nmatches and a checksum of vnMatches12.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "Features/ORBmatcher.hpp"

using namespace ORB_SLAM_Tracking;

int main(int argc, char** argv) {
  if (argc < 5) return 2;
  const int W = atoi(argv[2]), H = atoi(argv[3]), nf = atoi(argv[4]);
  std::vector<unsigned char> buf((size_t)2 * W * H);
  FILE* f = fopen(argv[1], "rb");
  if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) return 3;
  fclose(f);
  cv::Mat im1(H, W, CV_8UC1, buf.data()), im2(H, W, CV_8UC1, buf.data() + (size_t)W * H);
  float k[9] = {609.2855f, 0, 351.4274f, 0, 609.3422f, 237.7324f, 0, 0, 1};
  float d[4] = {-0.3492f, 0.1363f, 0, 0};
  cv::Mat K(3, 3, CV_32F, k), D(4, 1, CV_32F, d);
  ORBextractor ex(nf, 1.2f, 8, 20, 7);
  Frame F1(im1, 0.0, &ex, nullptr, K, D), F2(im2, 1.0, &ex, nullptr, K, D);
  std::vector<int> m12;
  ORBmatcher matcher(0.9f, true);
  const int n = matcher.SearchForInitialization(F1, F2, m12, 100);
  unsigned long long sum = 0;
  for (size_t i = 0; i < m12.size(); ++i) sum = sum * 31ull + (unsigned long long)(m12[i] + 1);
  std::printf("RESULT %d %d %d %llu %d\n", F1.N, F2.N, n, sum, ORBmatcher::DescriptorDistance(F1.mDescriptors.row(0), F2.mDescriptors.row(0)));
  std::vector<size_t> a = F2.GetFeaturesInArea(300.f, 200.f, 50.f, 0, 0);
  // public members and statics of the reference's classes
  std::printf("AREA %zu levels %d pyr %zu l1 %dx%d\n", a.size(), ex.GetLevels(), ex.mvImagePyramid.size(), ex.mvImagePyramid[1].cols,
              ex.mvImagePyramid[1].rows);
  std::printf("STATICS %.4f %.4f %d %d %d %d next %lu ids %lu %lu\n", Frame::fx, Frame::cy, Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY,
              Frame::mnMaxY, Frame::nNextId, F1.mnId, F2.mnId);
  Frame F3(F2);  // copy constructor: deep copy, same grid
  F3.SetPose(PoseT());
  std::printf("COPY %d %d %zu %d\n", F3.N, (int)(F3.mDescriptors.data != F2.mDescriptors.data), F3.mGrid[30][20].size() + 1 - F2.mGrid[30][20].size(),
              (int)F3.mbPoseSet);
  ExtractorNode node, c1, c2, c3, c4;  // Features/ORBextractor.hpp:32-53
  node.UL = cv::Point2i(0, 0); node.UR = cv::Point2i(101, 0); node.BL = cv::Point2i(0, 51); node.BR = cv::Point2i(101, 51);
  for (int i = 0; i < 5; ++i) { cv::KeyPoint k; k.pt.x = 25.f * i; k.pt.y = 12.f * i; node.vKeys.push_back(k); }
  node.DivideNode(c1, c2, c3, c4);
  std::printf("NODE %d %d %zu %zu %zu %zu %d\n", c1.UR.x, c1.BL.y, c1.vKeys.size(), c2.vKeys.size(), c3.vKeys.size(), c4.vKeys.size(), (int)c4.bNoMore);
  // single-frame latency through the reference's own call shape: Frame::Frame per image, SearchForInitialization per pair
  ex.SetFillImagePyramid(false);
  std::vector<double> tf, tm;
  for (int it = 0; it < 60; ++it) {
    auto t0 = std::chrono::steady_clock::now();
    Frame G1(it & 1 ? im2 : im1, 2.0 + it, &ex, nullptr, K, D);
    auto t1 = std::chrono::steady_clock::now();
    FILE* so = stdout; (void)so;
    std::vector<int> mm;
    auto t2 = std::chrono::steady_clock::now();
    matcher.SearchForInitialization(F1, G1, mm, 100);
    auto t3 = std::chrono::steady_clock::now();
    if (it >= 10) { tf.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count()); tm.push_back(std::chrono::duration<double, std::micro>(t3 - t2).count()); }
  }
  std::sort(tf.begin(), tf.end()); std::sort(tm.begin(), tm.end());
  std::printf("LATENCY frame_us p50 %.1f p99 %.1f match_us p50 %.1f p99 %.1f\n", tf[tf.size() / 2], tf[tf.size() - 1], tm[tm.size() / 2], tm[tm.size() - 1]);
  return 0;
}
