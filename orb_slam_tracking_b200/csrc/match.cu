// ORBmatcher::SearchForInitialization (Features/ORBmatcher.cpp:11-150), ComputeThreeMaxima (:152-183),
// Frame::GetFeaturesInArea (SlamTypes/Frame.cpp:163-206) and DBoW2::FORB::distance
// (Thirdparty/DBoW2/src/FORB.cpp:77-100) for a batch of frame pairs: one warp per pair.
//
// The reference loop over i1 is order-dependent (`vMatchedDistance[i2] <= dist` skips a candidate
// that an EARLIER query already matched at least as well, :67), so queries are walked in
// ascending i1 by the warp while the candidates of a query are spread over the lanes.
// Candidates are visited in the reference's order (grid column ix, then row iy, then insertion
// order), which for a fixed ix is one contiguous CSR range; the best/second-best reduction
// breaks distance ties by that position (strict `<` keeps the first minimum, :70-76).
// Descriptors are read as two 128-bit loads and compared with 8 x __popc.
#include <limits.h>

#include "common.cuh"

namespace orbx {

constexpr int TH_LOW = 50, HISTO_LENGTH = 30;

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const u8* __restrict__ d) {
  const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(d));
  const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(d) + 1);
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// One warp per pair.  Per query: the <= 64 grid columns of the search window are spread over the
// lanes (their CSR ranges are loaded in parallel and prefix-summed), then the flattened candidate
// list is walked 32 at a time; each lane keeps its two smallest distances and the visiting position
// of its smallest, and two redux.sync min-reductions give the reference's best / second best.
__global__ void __launch_bounds__(32) k_match(const __grid_constant__ Params P, const MatchArgs A, int pairStep,
                                              float nnratio, int checkOri, int window) {
  const int pair = blockIdx.x, lane = threadIdx.x;
  const int s1 = pair * pairStep, s2 = s1 + 1;
  const int n1 = A.n[s1], n2 = A.n[s2];
  const orbx_keypoint* __restrict__ k1 = A.keysUn + (size_t)s1 * A.stride;
  const orbx_keypoint* __restrict__ k2 = A.keysUn + (size_t)s2 * A.stride;
  const u8* __restrict__ d1 = A.desc + (size_t)s1 * A.stride * 32;
  const u8* __restrict__ d2 = A.desc + (size_t)s2 * A.stride * 32;
  const u32* __restrict__ gOff = A.gridOff + (size_t)s2 * (kGridCells + 1);
  const u32* __restrict__ gIdx = A.gridIdx + (size_t)s2 * A.stride;
  int* m12 = A.matches12 + (size_t)pair * A.stride;
  int* matchedDist = A.scratch + (size_t)pair * 3 * A.stride;
  int* m21 = matchedDist + A.stride;
  int* binOf = m21 + A.stride;
  __shared__ int sHist[HISTO_LENGTH];
  __shared__ int sStart[kGridCols + 1];  // exclusive prefix of the column range sizes
  __shared__ int sBeg[kGridCols];        // CSR start of each column range

  for (int i = lane; i < n1; i += 32) { m12[i] = -1; binOf[i] = -1; }
  for (int i = lane; i < n2; i += 32) { matchedDist[i] = INT_MAX; m21[i] = -1; }
  if (lane < HISTO_LENGTH) sHist[lane] = 0;
  __syncwarp();

  const float r = (float)window;
  const float minX = (float)P.bounds[0], minY = (float)P.bounds[2];
  const float factor = (float)HISTO_LENGTH / 360.0f;
  int nmatches = 0, badDist = 0, badRatio = 0;
  constexpr u32 kNone = 0xffffffffu;

  for (int i1 = 0; i1 < n1; ++i1) {
    const orbx_keypoint kp1 = k1[i1];
    if (kp1.octave > 0) continue;  // :38-39
    const float x = kp1.x, y = kp1.y;
    // GetFeaturesInArea(x, y, r, level1, level1)
    const int minCX = max(0, (int)floorf(__fmul_rn(x - minX - r, P.gridWInv)));
    if (minCX >= kGridCols) continue;
    const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(x - minX + r, P.gridWInv)));
    if (maxCX < 0) continue;
    const int minCY = max(0, (int)floorf(__fmul_rn(y - minY - r, P.gridHInv)));
    if (minCY >= kGridRows) continue;
    const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(y - minY + r, P.gridHInv)));
    if (maxCY < 0) continue;
    const int nCols = maxCX - minCX + 1;
    if (nCols <= 0) continue;
    // column ranges, two per lane
    int b0 = 0, c0 = 0, b1 = 0, c1 = 0;
    if (lane < nCols) {
      const int cell = (minCX + lane) * kGridRows;
      b0 = (int)gOff[cell + minCY];
      c0 = (int)gOff[cell + maxCY + 1] - b0;
    }
    if (lane + 32 < nCols) {
      const int cell = (minCX + lane + 32) * kGridRows;
      b1 = (int)gOff[cell + minCY];
      c1 = (int)gOff[cell + maxCY + 1] - b1;
    }
    int p0 = c0, p1 = c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t0 = __shfl_up_sync(0xffffffffu, p0, o), t1 = __shfl_up_sync(0xffffffffu, p1, o);
      if (lane >= o) { p0 += t0; p1 += t1; }
    }
    const int tot0 = __shfl_sync(0xffffffffu, p0, 31);
    const int total = tot0 + __shfl_sync(0xffffffffu, p1, 31);
    if (total == 0) continue;  // vIndices2.empty(), :46
    sStart[lane] = p0 - c0; sBeg[lane] = b0;
    sStart[lane + 32] = tot0 + p1 - c1; sBeg[lane + 32] = b1;
    if (lane == 0) sStart[kGridCols] = total;
    __syncwarp();
    const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32));
    const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32) + 1);
    u32 key = kNone;          // (best distance << 20) | visiting position
    int bidx = -1, bd2 = INT_MAX, found = 0;
    for (int t = lane; t < total; t += 32) {
      // column of flattened position t: last c with sStart[c] <= t
      int lo = 0, hi = kGridCols;
#pragma unroll
      for (int s = 0; s < 6; ++s) {
        const int mid = (lo + hi) >> 1;
        if (sStart[mid] <= t) lo = mid; else hi = mid;
      }
      const int i2 = (int)gIdx[sBeg[lo] + (t - sStart[lo])];
      const orbx_keypoint kp2 = k2[i2];
      if (kp2.octave != kp1.octave) continue;  // level filter [level1, level1]
      const float dx = kp2.x - x, dy = kp2.y - y;
      if (!(fabsf(dx) < r && fabsf(dy) < r)) continue;
      found = 1;
      const int dist = hamming256(a0, a1, d2 + (size_t)i2 * 32);
      if (matchedDist[i2] <= dist) continue;  // :67
      const u32 k = ((u32)dist << 20) | (u32)t;
      if (k < key) {  // lane-local positions ascend, so "dist < best" and "k < key" agree
        if (key != kNone) bd2 = (int)(key >> 20);
        key = k; bidx = i2;
      } else if (dist < bd2) bd2 = dist;
    }
    __syncwarp();
    if (!__any_sync(0xffffffffu, found)) continue;  // no candidate passed the window test
    const u32 best = __reduce_min_sync(0xffffffffu, key);
    const int winner = __ffs((int)__ballot_sync(0xffffffffu, key == best)) - 1;
    const int bestDist = best == kNone ? INT_MAX : (int)(best >> 20);
    const int mine2 = (lane == winner) ? bd2 : (key == kNone ? INT_MAX : (int)(key >> 20));
    const int bestDist2 = (int)__reduce_min_sync(0xffffffffu, (u32)mine2);
    const int bestIdx = __shfl_sync(0xffffffffu, bidx, winner);
    if (bestDist > TH_LOW) { ++badDist; continue; }                                  // :82
    if ((float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) { ++badRatio; continue; }  // :87
    const int prev = m21[bestIdx];
    if (prev >= 0) --nmatches;  // :95-98
    ++nmatches;
    int bin = -1;
    if (checkOri) {
      float rot = kp1.angle - k2[bestIdx].angle;
      if (rot < 0.0f) rot = rot + 360.0f;
      bin = (int)roundf(__fmul_rn(rot, factor));
      if (bin == HISTO_LENGTH) bin = 0;
    }
    __syncwarp();
    if (lane == 0) {
      if (prev >= 0) m12[prev] = -1;
      m12[i1] = bestIdx;
      m21[bestIdx] = i1;
      matchedDist[bestIdx] = bestDist;
      if (bin >= 0) { binOf[i1] = bin; sHist[bin] += 1; }
    }
    __syncwarp();
  }

  int badOri = 0;
  if (checkOri) {
    __syncwarp();
    // ComputeThreeMaxima, :152-183
    int i1m = -1, i2m = -1, i3m = -1, m1 = 0, m2 = 0, m3 = 0;
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      const int s = sHist[i];
      if (s > m1) { m3 = m2; m2 = m1; m1 = s; i3m = i2m; i2m = i1m; i1m = i; }
      else if (s > m2) { m3 = m2; m2 = s; i3m = i2m; i2m = i; }
      else if (s > m3) { m3 = s; i3m = i; }
    }
    if ((float)m2 < __fmul_rn(0.1f, (float)m1)) { i2m = -1; i3m = -1; }
    else if ((float)m3 < __fmul_rn(0.1f, (float)m1)) { i3m = -1; }
    for (int i = lane; i < n1; i += 32) {
      const int bn = binOf[i];
      if (bn >= 0 && bn != i1m && bn != i2m && bn != i3m) { m12[i] = -1; ++badOri; }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) badOri += __shfl_xor_sync(0xffffffffu, badOri, o);
    nmatches -= badOri;
  }
  if (lane == 0) {
    A.nmatches[pair] = nmatches;
    A.counters[pair * 3 + 0] = badDist;
    A.counters[pair * 3 + 1] = badRatio;
    A.counters[pair * 3 + 2] = badOri;
  }
}

int launch_match(const Params& P, const MatchArgs& A, int nPairs, int pairStep, float nnratio, int checkOri,
                 int window, cudaStream_t st) {
  if (nPairs <= 0) return 0;
  k_match<<<nPairs, 32, 0, st>>>(P, A, pairStep, nnratio, checkOri, window);
  return 1;
}

// n independent 256-bit Hamming distances (DescriptorDistance / FORB::distance)
__global__ void k_hamming(const u8* __restrict__ a, const u8* __restrict__ b, int n, int* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(a + (size_t)i * 32));
  const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(a + (size_t)i * 32) + 1);
  out[i] = hamming256(a0, a1, b + (size_t)i * 32);
}
void launch_hamming(const u8* a, const u8* b, int n, int* out, cudaStream_t st) {
  if (n > 0) k_hamming<<<(n + 255) / 256, 256, 0, st>>>(a, b, n, out);
}

// Frame::GetFeaturesInArea (SlamTypes/Frame.cpp:163-206) as a standalone query: one warp, ordered
// compaction with ballots so that the output keeps the reference's visiting order.
__global__ void __launch_bounds__(32) k_features_in_area(const __grid_constant__ Params P, const orbx_keypoint* __restrict__ kun,
                                                         const u32* __restrict__ gOff, const u32* __restrict__ gIdx,
                                                         float x, float y, float r, int minLevel, int maxLevel,
                                                         int* out, int cap, int* count) {
  const int lane = threadIdx.x;
  const float minX = (float)P.bounds[0], minY = (float)P.bounds[2];
  int n = 0;
  const int minCX = max(0, (int)floorf(__fmul_rn(x - minX - r, P.gridWInv)));
  const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(x - minX + r, P.gridWInv)));
  const int minCY = max(0, (int)floorf(__fmul_rn(y - minY - r, P.gridHInv)));
  const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(y - minY + r, P.gridHInv)));
  if (minCX < kGridCols && maxCX >= 0 && minCY < kGridRows && maxCY >= 0) {
    const bool check = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = minCX; ix <= maxCX; ++ix) {
      const int beg = (int)gOff[ix * kGridRows + minCY], end = (int)gOff[ix * kGridRows + maxCY + 1];
      for (int t0 = beg; t0 < end; t0 += 32) {
        const int t = t0 + lane;
        bool ok = false;
        int i2 = -1;
        if (t < end) {
          i2 = (int)gIdx[t];
          const orbx_keypoint k = kun[i2];
          ok = !(check && !(k.octave >= minLevel && k.octave <= maxLevel));
          ok = ok && fabsf(k.x - x) < r && fabsf(k.y - y) < r;
        }
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (ok) {
          const int p = n + __popc(m & ((1u << lane) - 1));
          if (p < cap) out[p] = i2;
        }
        n += __popc(m);
      }
    }
  }
  if (lane == 0) *count = n;
}
void launch_features_in_area(const Params& P, const orbx_keypoint* kun, const u32* gOff, const u32* gIdx, float x,
                             float y, float r, int minLevel, int maxLevel, int* out, int cap, int* count,
                             cudaStream_t st) {
  k_features_in_area<<<1, 32, 0, st>>>(P, kun, gOff, gIdx, x, y, r, minLevel, maxLevel, out, cap, count);
}

}  // namespace orbx
