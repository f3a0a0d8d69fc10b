// Shared pieces of the matcher kernels (match.cu, match_bf.cu).
#pragma once
#include "common.cuh"

namespace orbx {

constexpr int TH_LOW = 50, HISTO_LENGTH = 30;  // Features/ORBmatcher.cpp:5-7
constexpr int kBfK = 7;                        // list length of k_bf_topk (a row is kBfK keys + one info word = 32 bytes)

// DBoW2::FORB::distance (Thirdparty/DBoW2/src/FORB.cpp:77-100): 256-bit popcount of the XOR
__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const u8* __restrict__ d) {
  const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(d));
  const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(d) + 1);
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// the window's cell rectangle, Frame::GetFeaturesInArea :167-185; false = empty result
__device__ __forceinline__ bool cell_rect(const Params& P, float x, float y, float r, int& minCX, int& maxCX, int& minCY,
                                          int& maxCY) {
  const float minX = (float)P.bounds[0], minY = (float)P.bounds[2];
  minCX = max(0, (int)floorf(__fmul_rn(x - minX - r, P.gridWInv)));
  if (minCX >= kGridCols) return false;
  maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(x - minX + r, P.gridWInv)));
  if (maxCX < 0) return false;
  minCY = max(0, (int)floorf(__fmul_rn(y - minY - r, P.gridHInv)));
  if (minCY >= kGridRows) return false;
  maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(y - minY + r, P.gridHInv)));
  if (maxCY < 0) return false;
  return maxCX >= minCX && maxCY >= minCY;
}

// ORBmatcher::ComputeThreeMaxima, Features/ORBmatcher.cpp:152-183 (on the histogram's bin counts)
__device__ __forceinline__ void three_maxima(const int* hist, int& i1m, int& i2m, int& i3m) {
  int m1 = 0, m2 = 0, m3 = 0;
  i1m = i2m = i3m = -1;
  for (int i = 0; i < HISTO_LENGTH; ++i) {
    const int h = hist[i];
    if (h > m1) { m3 = m2; m2 = m1; m1 = h; i3m = i2m; i2m = i1m; i1m = i; }
    else if (h > m2) { m3 = m2; m2 = h; i3m = i2m; i2m = i; }
    else if (h > m3) { m3 = h; i3m = i; }
  }
  if ((float)m2 < __fmul_rn(0.1f, (float)m1)) { i2m = -1; i3m = -1; }
  else if ((float)m3 < __fmul_rn(0.1f, (float)m1)) { i3m = -1; }
}

}  // namespace orbx
