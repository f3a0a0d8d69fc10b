// Host check of stdsort_emul() against the real std::sort of this libstdc++ (the reference's
// DistributeOctTree uses std::sort with a comparator that ties, Features/ORBextractor.cpp:684-696,912).
// Prints the number of mismatching trials; exit code 0 iff none.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <utility>
#include <vector>

#include "../../orb_slam_tracking_b200/csrc/stdsort_emul.cuh"

struct Node { int ulx; };
typedef std::pair<int, Node*> SN;
static bool cmp(SN& l, SN& r) {  // same shape as the reference's compareNodes (non-const refs)
  if (l.first < r.first) return true;
  if (l.first > r.first) return false;
  return l.second->ulx < r.second->ulx;
}

int main(int argc, char** argv) {
  const int trials = argc > 1 ? atoi(argv[1]) : 3000;
  std::mt19937 rng(12345);
  long bad = 0;
  for (int t = 0; t < trials; ++t) {
    int n;
    const int mode = t % 6;
    if (mode == 0) n = rng() % 17;
    else if (mode == 1) n = 17 + rng() % 64;
    else if (mode == 2) n = 1 + rng() % 400;
    else if (mode == 3) n = 1 + rng() % 3000;
    else if (mode == 4) n = 1 + rng() % 20000;
    else n = 16 + rng() % 3;
    const int kc = 1 + rng() % (mode == 4 ? 50 : 6);   // few distinct counts -> many ties
    const int kx = 1 + rng() % 8;
    std::vector<Node> nodes(n);
    std::vector<SN> ref(n);
    std::vector<orbx::SortElem> emu(n);
    for (int i = 0; i < n; ++i) {
      int c = 2 + rng() % kc, x = (rng() % kx) * 37;
      if (t % 11 == 0) { c = 2 + i % kc; }            // structured inputs
      if (t % 13 == 0) { c = 2 + (n - i) % kc; }
      if (t % 17 == 0) { c = 5; x = 0; }              // all equal
      nodes[i].ulx = x;
      ref[i] = SN(c, &nodes[i]);
      emu[i].key = ((unsigned long long)(unsigned)c << 32) | (unsigned)x;
      emu[i].id = (unsigned)i; emu[i].pad = 0;
    }
    std::sort(ref.begin(), ref.end(), cmp);
    orbx::stdsort_emul(emu.data(), n);
    for (int i = 0; i < n; ++i)
      if ((int)(ref[i].second - nodes.data()) != (int)emu[i].id) { ++bad; break; }
  }
  // killer sequence for median-of-3 quicksort to exercise the heap-sort fallback
  {
    const int n = 4096;
    std::vector<Node> nodes(n);
    std::vector<SN> ref(n);
    std::vector<orbx::SortElem> emu(n);
    std::vector<int> v(n);
    // organ-pipe / anti-quicksort-ish patterns
    for (int pat = 0; pat < 3; ++pat) {
      for (int i = 0; i < n; ++i) {
        int c = pat == 0 ? (i < n / 2 ? i : n - i) : pat == 1 ? ((i & 1) ? i : n - i) : (i * 7919) % 13;
        nodes[i].ulx = (i * 31) % 5;
        ref[i] = SN(c, &nodes[i]);
        emu[i].key = ((unsigned long long)(unsigned)c << 32) | (unsigned)nodes[i].ulx;
        emu[i].id = (unsigned)i; emu[i].pad = 0;
      }
      std::sort(ref.begin(), ref.end(), cmp);
      orbx::stdsort_emul(emu.data(), n);
      for (int i = 0; i < n; ++i)
        if ((int)(ref[i].second - nodes.data()) != (int)emu[i].id) { ++bad; break; }
    }
  }

  // McIlroy's "antiqsort" adversary builds an input on which this libstdc++'s median-of-3
  // quicksort degenerates, so the introsort depth limit trips and the heap-sort fallback runs.
  {
    const int n = 6000;
    static std::vector<int> val; static int nsolid, candidate, gas;
    val.assign(n, 0); gas = n; nsolid = 0; candidate = 0;
    for (int i = 0; i < n; ++i) val[i] = gas;
    std::vector<int> idx(n);
    for (int i = 0; i < n; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [](int x, int y) {
      if (val[x] == gas && val[y] == gas) { if (x == candidate) val[x] = nsolid++; else val[y] = nsolid++; }
      if (val[x] == gas) candidate = x; else if (val[y] == gas) candidate = y;
      return val[x] < val[y];
    });
    for (int q = 1; q <= 8; q *= 2) {   // q > 1 quantises the keys -> ties inside the killer input
      std::vector<Node> nodes(n);
      std::vector<SN> ref(n);
      std::vector<orbx::SortElem> emu(n);
      for (int i = 0; i < n; ++i) {
        nodes[i].ulx = 0;
        ref[i] = SN(val[i] / q, &nodes[i]);
        emu[i].key = ((unsigned long long)(unsigned)(val[i] / q) << 32);
        emu[i].id = (unsigned)i; emu[i].pad = 0;
      }
      std::sort(ref.begin(), ref.end(), cmp);
      orbx::stdsort_emul(emu.data(), n);
      for (int i = 0; i < n; ++i)
        if ((int)(ref[i].second - nodes.data()) != (int)emu[i].id) { ++bad; break; }
    }
  }
  printf("%ld\n", bad);
  return bad ? 1 : 0;
}
