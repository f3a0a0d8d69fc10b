// Host check of stdnth_emul_t() against the real std::nth_element of this libstdc++, with the comparator
// cv::KeyPointsFilter::retainBest uses (response greater) on keypoint-like records with many tied responses
// (ComputeKeyPointsOld, Features/ORBextractor.cpp:1452,1482).  Prints the number of mismatching trials; exit code 0 iff none.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../../orb_slam_tracking_b200/csrc/stdsort_emul.cuh"

struct KP { float x, y, size, angle, response; int octave, class_id; };
struct Greater { bool operator()(const KP& a, const KP& b) const { return a.response > b.response; } };
struct E { unsigned long long w; };   // the device record: response in the high bits, payload below
struct EGreater { bool operator()(const E& a, const E& b) const { return (a.w >> 48) > (b.w >> 48); } };

int main(int argc, char** argv) {
  const int trials = argc > 1 ? atoi(argv[1]) : 4000;
  std::mt19937 rng(4242);
  long bad = 0;
  for (int t = 0; t < trials; ++t) {
    int n;
    const int mode = t % 5;
    if (mode == 0) n = 1 + rng() % 8;
    else if (mode == 1) n = 1 + rng() % 64;
    else if (mode == 2) n = 1 + rng() % 600;
    else if (mode == 3) n = 1 + rng() % 5000;
    else n = 2 + rng() % 3;
    const int kr = 1 + rng() % (mode == 3 ? 200 : 12);  // few distinct responses: FAST scores are small integers
    std::vector<KP> ref(n);
    std::vector<E> emu(n);
    for (int i = 0; i < n; ++i) {
      int r = 8 + rng() % kr;
      if (t % 7 == 0) r = 8 + i % kr;
      if (t % 11 == 0) r = 8 + (n - i) % kr;
      if (t % 13 == 0) r = 20;
      ref[i] = KP{(float)i, 0, 7, -1, (float)r, 0, -1};
      emu[i].w = ((unsigned long long)r << 48) | (unsigned long long)i;
    }
    const int nth = rng() % n;
    std::nth_element(ref.begin(), ref.begin() + nth, ref.end(), Greater());
    orbx::stdnth_emul_t(emu.data(), n, nth, EGreater());
    for (int i = 0; i < n; ++i)
      if ((int)ref[i].x != (int)(emu[i].w & 0xffffffffull)) { ++bad; break; }
  }
  // adversary for median-of-3 selection: drives __introselect into its heap-select branch
  for (int rep = 0; rep < 4; ++rep) {
    const int n = 6000;
    std::vector<int> v(n);
    {  // anti-quicksort (McIlroy) against std::nth_element
      std::vector<int> val(n, n), idx(n);
      for (int i = 0; i < n; ++i) idx[i] = i;
      int nsolid = 0, candidate = 0;
      auto cmp = [&](int a, int b) {
        if (val[a] == n && val[b] == n) { if (a == candidate) val[a] = nsolid++; else val[b] = nsolid++; }
        if (val[a] == n) candidate = a; else if (val[b] == n) candidate = b;
        return val[a] < val[b];
      };
      std::nth_element(idx.begin(), idx.begin() + (rep * n) / 5, idx.end(), cmp);
      for (int i = 0; i < n; ++i) v[i] = val[i] == n ? nsolid++ : val[i];
    }
    std::vector<KP> ref(n);
    std::vector<E> emu(n);
    for (int i = 0; i < n; ++i) {
      ref[i] = KP{(float)i, 0, 7, -1, (float)(n - v[i]), 0, -1};   // comparator is "greater": mirror the values
      emu[i].w = ((unsigned long long)(n - v[i]) << 48) | (unsigned long long)i;
    }
    const int nth = (rep * n) / 5;
    std::nth_element(ref.begin(), ref.begin() + nth, ref.end(), Greater());
    orbx::stdnth_emul_t(emu.data(), n, nth, EGreater());
    for (int i = 0; i < n; ++i)
      if ((int)ref[i].x != (int)(emu[i].w & 0xffffffffull)) { ++bad; break; }
  }
  printf("%ld\n", bad);
  return bad ? 1 : 0;
}
