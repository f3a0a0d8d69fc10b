// Exercises the C++ drop-in headers the way demo_initialization.cpp:65-108 uses the reference classes:
// two gray images -> ORBextractor -> two Frames -> ORBmatcher(0.9, true).SearchForInitialization.
// Input: raw u8 file with two W x H images.  Prints N1 N2 nmatches and a checksum of vnMatches12.
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
  std::printf("AREA %zu levels %d pyr %zu\n", a.size(), ex.GetLevels(), ex.FetchImagePyramid().size());
  return 0;
}
