// ORBmatcher::SearchForInitialization (Features/ORBmatcher.cpp:11-150), ComputeThreeMaxima (:152-183),
// Frame::GetFeaturesInArea (SlamTypes/Frame.cpp:163-206) and DBoW2::FORB::distance
// (Thirdparty/DBoW2/src/FORB.cpp:77-100) for a batch of frame pairs.  k_match_cta (below) is the product path;
// k_match (one warp per pair, everything in one ordered walk) remains for keypoint capacities whose tables do
// not fit shared memory.
//
// The reference loop over i1 is order-dependent (`vMatchedDistance[i2] <= dist` skips a candidate
// that an EARLIER query already matched at least as well, :67), so queries are walked in
// ascending i1 by the warp while the candidates of a query are spread over the lanes.
// Candidates are visited in the reference's order (grid column ix, then row iy, then insertion
// order), which for a fixed ix is one contiguous CSR range; the best/second-best reduction
// breaks distance ties by that position (strict `<` keeps the first minimum, :70-76).
// Descriptors are read as two 128-bit loads and compared with 8 x __popc.
#include <limits.h>

#include <algorithm>

#include <cstdlib>

#include "match_common.cuh"

namespace orbx {

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

// ---- the batched matcher: one CTA of 8 warps per pair, distances in parallel, resolve in order -------------
// The order dependence of the reference loop sits only in `vMatchedDistance[i2] <= dist` (:67) and in
// the bookkeeping after a match; which candidates a query sees, in which order, and their distances do
// not depend on earlier queries.  So per chunk of queries all warps fill (distance, i2) entries in the
// reference's visiting order into shared memory, and then ONE warp replays the reference's loop over
// those entries (ascending i1), with vMatchedDistance / vnMatches21 in shared memory.
//   S1  level-0 CSR of F2's grid (SearchForInitialization only asks for level [0, 0], :38-44): same cell
//       order and in-cell order as mGrid, other octaves dropped, candidate x / y / index staged beside it;
//   S2  the level-0 queries of F1 in ascending i1;
//   S3  per chunk: segment sizes (upper bound = level-0 keys in the window's cell rectangle), prefix sums,
//       warp-per-query fill (window test, 256-bit Hamming), then the ordered replay by warp 0.
constexpr int MT = 256, MW = MT / 32;
constexpr int kChunkQ = 128;            // queries per chunk (segment table size)
constexpr u32 kInvalid = 0xffffffffu;   // entry of a cell-rectangle key outside the window

struct MatchSmem {  // byte offsets into dynamic shared memory
  int off0, cx, cy, ci, md, m21, q, bin, touch, ent, d2, qd, qk, total;
  int entCap, descCap;
};
constexpr int kDescStage = 1024;  // level-0 keys of F2 whose descriptors are staged in shared memory (the rest is read from global memory)
__host__ __device__ inline MatchSmem match_smem_layout(int stride) {
  MatchSmem L;
  int o = 0;
  auto take = [&](int bytes) { const int r = o; o += (bytes + 15) & ~15; return r; };
  L.off0 = take((kGridCells + 1) * 2);
  L.cx = take(stride * 4);
  L.cy = take(stride * 4);
  L.ci = take(stride * 2);
  L.md = take(stride * 2);
  L.m21 = take(stride * 2);
  L.q = take(stride * 2);
  L.bin = take(stride);
  L.touch = take(stride * 4);
  L.entCap = stride > 8192 ? stride : 8192;
  L.ent = take(L.entCap * 4);
  // descriptors next to the work: F2's level-0 keys in CSR order (the first descCap of them) and the queries of the running chunk.
  // Without them a warp's fill is a chain of dependent global loads per query (keypoint, its descriptor, the candidates' descriptors).
  L.descCap = stride < kDescStage ? stride : kDescStage;
  L.d2 = take(L.descCap * 32);
  L.qd = take(kChunkQ * 32);
  L.qk = take(kChunkQ * 8);
  L.total = o;
  return L;
}

__global__ void __launch_bounds__(MT) k_match_cta(const __grid_constant__ Params P, const MatchArgs A, int pairStep, float nnratio,
                                                  int checkOri, int window) {
  extern __shared__ __align__(16) u8 dsm[];
  __shared__ int sPart[MT];
  __shared__ int sSeg[kChunkQ + 1];
  __shared__ int sColStart[MW][kGridCols + 1];
  __shared__ int sColBeg[MW][kGridCols];
  __shared__ int sHist[HISTO_LENGTH];
  __shared__ int sNQ, sChunkN;
  const MatchSmem LY = match_smem_layout(A.stride);
  unsigned short* const sOff0 = reinterpret_cast<unsigned short*>(dsm + LY.off0);
  float* const sCx = reinterpret_cast<float*>(dsm + LY.cx);
  float* const sCy = reinterpret_cast<float*>(dsm + LY.cy);
  unsigned short* const sCi = reinterpret_cast<unsigned short*>(dsm + LY.ci);
  unsigned short* const sMD = reinterpret_cast<unsigned short*>(dsm + LY.md);
  unsigned short* const sM21 = reinterpret_cast<unsigned short*>(dsm + LY.m21);
  unsigned short* const sQ = reinterpret_cast<unsigned short*>(dsm + LY.q);
  u8* const sBin = dsm + LY.bin;
  u32* const sTouch = reinterpret_cast<u32*>(dsm + LY.touch);  // per key of F2: the first query (lane) of the running batch that takes it
  u32* const sEnt = reinterpret_cast<u32*>(dsm + LY.ent);
  uint4* const sD2 = reinterpret_cast<uint4*>(dsm + LY.d2);   // [p][2]: descriptor of the level-0 key at CSR position p < descCap
  uint4* const sQd = reinterpret_cast<uint4*>(dsm + LY.qd);   // [query of the chunk][2]
  float2* const sQk = reinterpret_cast<float2*>(dsm + LY.qk); // [query of the chunk]: undistorted position

  const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s1 = pair * pairStep, s2 = s1 + 1;
  const int n1 = A.n[s1], n2 = A.n[s2];
  const orbx_keypoint* __restrict__ k1 = A.keysUn + (size_t)s1 * A.stride;
  const orbx_keypoint* __restrict__ k2 = A.keysUn + (size_t)s2 * A.stride;
  const u8* __restrict__ d1 = A.desc + (size_t)s1 * A.stride * 32;
  const u8* __restrict__ d2 = A.desc + (size_t)s2 * A.stride * 32;
  const u32* __restrict__ gOff = A.gridOff + (size_t)s2 * (kGridCells + 1);
  const u32* __restrict__ gIdx = A.gridIdx + (size_t)s2 * A.stride;
  int* m12 = A.matches12 + (size_t)pair * A.stride;

  for (int i = tid; i < n1; i += MT) m12[i] = -1;
  for (int i = tid; i < n2; i += MT) { sMD[i] = 0xffffu; sM21[i] = 0xffffu; sTouch[i] = 0xffffffffu; }
  if (tid < HISTO_LENGTH) sHist[tid] = 0;

  // ---- S1: level-0 CSR of F2 (thread t owns cells [12 t, 12 t + 12)) ----
  // The grid entries of F2 (index -> keypoint: two dependent global loads each) are fetched by all threads side by side into a
  // scratch view of the entry table (x, y, index; index 0xffffffff for other octaves) when they fit there; the per-cell walks
  // below then read shared memory.  Otherwise they walk global memory as before.
  const int nE = (int)gOff[kGridCells];
  const bool staged = (size_t)nE * 12 <= (size_t)LY.entCap * 4;
  float* const tX = reinterpret_cast<float*>(sEnt);
  float* const tY = tX + nE;
  u32* const tI = reinterpret_cast<u32*>(tY + nE);
  if (staged) {
    for (int e = tid; e < nE; e += MT) {
      const int i2 = (int)gIdx[e];
      const orbx_keypoint k = k2[i2];
      tX[e] = k.x; tY[e] = k.y; tI[e] = k.octave == 0 ? (u32)i2 : 0xffffffffu;
    }
    __syncthreads();
  }
  constexpr int CPT = kGridCells / MT;
  int cnt[CPT], tot = 0;
  u32 eb[CPT + 1];
  {
    const int c0 = tid * CPT;
#pragma unroll
    for (int j = 0; j <= CPT; ++j) eb[j] = gOff[c0 + j];
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      int n = 0;
      if (staged) { for (u32 e = eb[j]; e < eb[j + 1]; ++e) n += (tI[e] != 0xffffffffu); }
      else { for (u32 e = eb[j]; e < eb[j + 1]; ++e) n += (k2[gIdx[e]].octave == 0); }
      cnt[j] = n;
      tot += n;
    }
  }
  int total0;
  int run0 = cta_prefix_sum<MT>(tot, sPart, &total0) - tot;
  {
    int run = run0;
    const int c0 = tid * CPT;
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      sOff0[c0 + j] = (unsigned short)run;
      for (u32 e = eb[j]; e < eb[j + 1]; ++e) {
        if (staged) {
          const u32 i2 = tI[e];
          if (i2 != 0xffffffffu) { sCx[run] = tX[e]; sCy[run] = tY[e]; sCi[run] = (unsigned short)i2; ++run; }
        } else {
          const int i2 = (int)gIdx[e];
          const orbx_keypoint k = k2[i2];
          if (k.octave == 0) { sCx[run] = k.x; sCy[run] = k.y; sCi[run] = (unsigned short)i2; ++run; }
        }
      }
    }
    if (tid == MT - 1) sOff0[kGridCells] = (unsigned short)run;
  }
  __syncthreads();
  {  // descriptors of F2's level-0 keys, CSR order (visible after the barriers of S2)
    const int n0 = min((int)sOff0[kGridCells], LY.descCap);
    for (int i = tid; i < 2 * n0; i += MT) sD2[i] = __ldg(reinterpret_cast<const uint4*>(d2 + (size_t)sCi[i >> 1] * 32) + (i & 1));
  }
  // ---- S2: level-0 queries of F1, ascending i1 (:36-39) ----
  {
    const int per = (n1 + MT - 1) / MT, b = min(tid * per, n1), e = min(b + per, n1);
    int c = 0;
    for (int i = b; i < e; ++i) c += !(k1[i].octave > 0);
    int totalQ;
    int run = cta_prefix_sum<MT>(c, sPart, &totalQ) - c;
    for (int i = b; i < e; ++i)
      if (!(k1[i].octave > 0)) { sQ[run] = (unsigned short)i; sBin[run] = 0xff; ++run; }
    if (tid == MT - 1) sNQ = run;
  }
  __syncthreads();
  const int nQ = sNQ;
  const float r = (float)window;
  const float factor = (float)HISTO_LENGTH / 360.0f;
  int nmatches = 0, badDist = 0, badRatio = 0;  // live in warp 0
  constexpr u32 kNone = 0xffffffffu;

  for (int qs = 0; qs < nQ;) {
    // (a) segment sizes, one thread per query of the chunk
    const int nCand = min(kChunkQ, nQ - qs);
    if (tid < kChunkQ) {
      int sz = 0;
      if (tid < nCand) {
        const int i1 = sQ[qs + tid];
        const orbx_keypoint kp1 = k1[i1];
        sQk[tid] = make_float2(kp1.x, kp1.y);
        sQd[2 * tid] = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32));
        sQd[2 * tid + 1] = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32) + 1);
        int a, b, c, d;
        if (cell_rect(P, kp1.x, kp1.y, r, a, b, c, d))
          for (int ix = a; ix <= b; ++ix) sz += (int)sOff0[ix * kGridRows + d + 1] - (int)sOff0[ix * kGridRows + c];
      }
      sPart[tid] = sz;
    }
    __syncthreads();
    // (b) exclusive prefix over the 128 sizes; the chunk takes the queries whose entries fit
    if (warp == 0) {
      int v[4], sum = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) { v[j] = sPart[lane * 4 + j]; sum += v[j]; }
      int inc = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      int start = inc - sum, fit = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sSeg[lane * 4 + j] = start;
        start += v[j];
        fit += (lane * 4 + j < nCand && start <= LY.entCap);
      }
      if (lane == 31) sSeg[kChunkQ] = start;
#pragma unroll
      for (int o = 16; o; o >>= 1) fit += __shfl_xor_sync(0xffffffffu, fit, o);
      if (lane == 0) sChunkN = fit;
    }
    __syncthreads();
    const int nq = sChunkN;  // >= 1: one segment never exceeds the level-0 keys of F2 <= entCap
    // (c) fill, one warp per query: flattened walk over the column ranges of the cell rectangle
    for (int qi = warp; qi < nq; qi += MW) {
      const int segB = sSeg[qi], segN = sSeg[qi + 1] - segB;
      if (segN == 0) continue;
      const float2 kp1 = sQk[qi];
      int minCX, maxCX, minCY, maxCY;
      cell_rect(P, kp1.x, kp1.y, r, minCX, maxCX, minCY, maxCY);
      const int nCols = maxCX - minCX + 1;
      int b0 = 0, c0 = 0, b1 = 0, c1 = 0;
      if (lane < nCols) {
        const int cell = (minCX + lane) * kGridRows;
        b0 = (int)sOff0[cell + minCY];
        c0 = (int)sOff0[cell + maxCY + 1] - b0;
      }
      if (lane + 32 < nCols) {
        const int cell = (minCX + lane + 32) * kGridRows;
        b1 = (int)sOff0[cell + minCY];
        c1 = (int)sOff0[cell + maxCY + 1] - b1;
      }
      int p0 = c0, p1 = c1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t0 = __shfl_up_sync(0xffffffffu, p0, o), t1 = __shfl_up_sync(0xffffffffu, p1, o);
        if (lane >= o) { p0 += t0; p1 += t1; }
      }
      const int tot0 = __shfl_sync(0xffffffffu, p0, 31);
      int* cs = sColStart[warp];
      int* cb = sColBeg[warp];
      cs[lane] = p0 - c0; cb[lane] = b0;
      cs[lane + 32] = tot0 + p1 - c1; cb[lane + 32] = b1;
      if (lane == 0) cs[kGridCols] = segN;
      __syncwarp();
      const uint4 a0 = sQd[2 * qi], a1 = sQd[2 * qi + 1];
      for (int t = lane; t < segN; t += 32) {
        int lo = 0, hi = kGridCols;  // column of flattened position t: last c with cs[c] <= t
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          const int mid = (lo + hi) >> 1;
          if (cs[mid] <= t) lo = mid; else hi = mid;
        }
        const int p = cb[lo] + (t - cs[lo]);
        const float dx = sCx[p] - kp1.x, dy = sCy[p] - kp1.y;
        u32 en = kInvalid;
        if (fabsf(dx) < r && fabsf(dy) < r) {  // Frame.cpp:196-199
          const int i2 = sCi[p];
          int dist;
          if (p < LY.descCap) {
            const uint4 b0 = sD2[2 * p], b1 = sD2[2 * p + 1];
            dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                   __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
          } else {
            dist = hamming256(a0, a1, d2 + (size_t)i2 * 32);
          }
          en = ((u32)dist << 16) | (u32)i2;
        }
        sEnt[segB + t] = en;
      }
      __syncwarp();
    }
    __syncthreads();
    // (d) the reference loop over the chunk's queries, in order (:36-117), by warp 0 -- 32 queries at a time.
    // A query's outcome depends on the earlier ones only through vMatchedDistance (:67): an earlier match to key c changes it
    // only if c is one of the query's two best unskipped candidates.  So every lane first decides ITS query against the
    // state at the start of the batch (its own candidates, in order); a lane whose best or second-best key is taken by an
    // earlier lane of the batch is in conflict.  All lanes before the first conflict commit together (their keys are
    // distinct); the conflicting query is then decided again against the current state by the whole warp (the reference's
    // loop body, below), and the rest of the batch is checked again.  The result is the ordered loop's, query for query.
    if (warp == 0) {
      // the reference's loop body for one query, lanes over its candidates; returns the matched key of F2 or -1
      auto resolve_one = [&](int qi) -> int {
        const int segB = sSeg[qi], segN = sSeg[qi + 1] - segB;
        if (segN == 0) return -1;  // vIndices2.empty(), :46
        u32 key = kNone;           // (best distance << 20) | visiting position
        int bidx = -1, bd2 = INT_MAX, found = 0;
        for (int t = lane; t < segN; t += 32) {
          const u32 en = sEnt[segB + t];
          if (en == kInvalid) continue;
          found = 1;
          const int dist = (int)(en >> 16), i2 = (int)(en & 0xffffu);
          if ((int)sMD[i2] <= dist) continue;  // :67
          const u32 k = ((u32)dist << 20) | (u32)t;
          if (k < key) {  // lane-local positions ascend, so "dist < best" and "k < key" agree
            if (key != kNone) bd2 = (int)(key >> 20);
            key = k; bidx = i2;
          } else if (dist < bd2) bd2 = dist;
        }
        if (!__any_sync(0xffffffffu, found)) return -1;  // no key of the rectangle is inside the window
        const u32 best = __reduce_min_sync(0xffffffffu, key);
        const int winner = __ffs((int)__ballot_sync(0xffffffffu, key == best)) - 1;
        const int bestDist = best == kNone ? INT_MAX : (int)(best >> 20);
        const int mine2 = (lane == winner) ? bd2 : (key == kNone ? INT_MAX : (int)(key >> 20));
        const int bestDist2 = (int)__reduce_min_sync(0xffffffffu, (u32)mine2);
        const int bestIdx = __shfl_sync(0xffffffffu, bidx, winner);
        if (bestDist > TH_LOW) { ++badDist; return -1; }                                        // :82
        if ((float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) { ++badRatio; return -1; }  // :87
        const int prev = sM21[bestIdx] == 0xffffu ? -1 : (int)sM21[bestIdx];
        if (prev >= 0) --nmatches;  // :95-98
        ++nmatches;
        const int i1 = sQ[qs + qi];
        __syncwarp();
        if (lane == 0) {
          if (prev >= 0) m12[prev] = -1;
          m12[i1] = bestIdx;
          sM21[bestIdx] = (unsigned short)i1;
          sMD[bestIdx] = (unsigned short)bestDist;
          if (checkOri) {
            float rot = k1[i1].angle - k2[bestIdx].angle;
            if (rot < 0.0f) rot = rot + 360.0f;
            int bin = (int)roundf(__fmul_rn(rot, factor));
            if (bin == HISTO_LENGTH) bin = 0;
            sBin[qs + qi] = (u8)bin;
            sHist[bin] += 1;
          }
        }
        __syncwarp();
        return bestIdx;
      };
      for (int qb = 0; qb < nq; qb += 32) {
        const int qi = qb + lane;
        int segB = 0, segN = 0;
        if (qi < nq) { segB = sSeg[qi]; segN = sSeg[qi + 1] - segB; }
        // every lane: its own query against the state at the start of the batch
        int best = INT_MAX, best2 = INT_MAX, bidx = -1, bidx2 = -1;
        bool found = false;
        const int maxN = (int)__reduce_max_sync(0xffffffffu, (u32)segN);
        for (int t = 0; t < maxN; ++t) {
          if (t < segN) {
            const u32 en = sEnt[segB + t];
            if (en != kInvalid) {
              found = true;
              const int dist = (int)(en >> 16), i2 = (int)(en & 0xffffu);
              if (!((int)sMD[i2] <= dist)) {  // :67
                if (dist < best) { best2 = best; bidx2 = bidx; best = dist; bidx = i2; }  // :70-76, strict <: the first minimum stays
                else if (dist < best2) { best2 = dist; bidx2 = i2; }
              }
            }
          }
        }
        const bool bad = best > TH_LOW;                                                    // :82 (also "every candidate skipped")
        const bool ratio = !bad && (float)best > __fmul_rn(nnratio, (float)best2);        // :87
        u32 pend = __ballot_sync(0xffffffffu, found);
        int myTouch = -1;
        while (pend) {
          const bool mine = (pend >> lane) & 1u;
          const bool isMatch = mine && !bad && !ratio;
          if (isMatch) atomicMin(&sTouch[bidx], (u32)lane);
          __syncwarp();
          // a query that fails on distance cannot be rescued by a skipped candidate (its best can only get worse)
          const bool conf = mine && !bad && (sTouch[bidx] < (u32)lane || (bidx2 >= 0 && sTouch[bidx2] < (u32)lane));
          const u32 cm = __ballot_sync(0xffffffffu, conf);
          const int f = cm ? __ffs((int)cm) - 1 : 32;
          const u32 commit = f >= 32 ? pend : pend & ((1u << f) - 1u);
          const bool doCommit = (commit >> lane) & 1u;
          __syncwarp();
          if (isMatch && !doCommit && sTouch[bidx] == (u32)lane) sTouch[bidx] = 0xffffffffu;  // tentative claims are withdrawn
          int prev = -1;
          if (doCommit && isMatch) {
            prev = sM21[bidx] == 0xffffu ? -1 : (int)sM21[bidx];
            const int i1 = sQ[qs + qi];
            if (prev >= 0) m12[prev] = -1;  // :95-98
            m12[i1] = bidx;
            sM21[bidx] = (unsigned short)i1;
            sMD[bidx] = (unsigned short)best;
            myTouch = bidx;
            if (checkOri) {
              float rot = k1[i1].angle - k2[bidx].angle;
              if (rot < 0.0f) rot = rot + 360.0f;
              int bin = (int)roundf(__fmul_rn(rot, factor));
              if (bin == HISTO_LENGTH) bin = 0;
              sBin[qs + qi] = (u8)bin;
              atomicAdd(&sHist[bin], 1);
            }
          }
          badDist += __popc(__ballot_sync(0xffffffffu, doCommit && bad));
          badRatio += __popc(__ballot_sync(0xffffffffu, doCommit && ratio));
          nmatches += __popc(__ballot_sync(0xffffffffu, doCommit && isMatch)) - __popc(__ballot_sync(0xffffffffu, prev >= 0));
          pend &= ~commit;
          __syncwarp();
          if (f < 32) {
            const int got = resolve_one(qb + f);
            if (got >= 0) {
              if (lane == 0) atomicMin(&sTouch[got], (u32)f);
              if (lane == f) myTouch = got;
            }
            pend &= ~(1u << f);
            __syncwarp();
          }
        }
        if (myTouch >= 0) sTouch[myTouch] = 0xffffffffu;
        __syncwarp();
      }
    }
    __syncthreads();
    qs += nq;
  }

  if (warp != 0) return;
  int badOri = 0;
  if (checkOri) {
    // ComputeThreeMaxima, :152-183
    int i1m = -1, i2m = -1, i3m = -1, m1 = 0, m2 = 0, m3 = 0;
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      const int h = sHist[i];
      if (h > m1) { m3 = m2; m2 = m1; m1 = h; i3m = i2m; i2m = i1m; i1m = i; }
      else if (h > m2) { m3 = m2; m2 = h; i3m = i2m; i2m = i; }
      else if (h > m3) { m3 = h; i3m = i; }
    }
    if ((float)m2 < __fmul_rn(0.1f, (float)m1)) { i2m = -1; i3m = -1; }
    else if ((float)m3 < __fmul_rn(0.1f, (float)m1)) { i3m = -1; }
    for (int qp = lane; qp < nQ; qp += 32) {
      const int bn = sBin[qp];
      if (bn != 0xff && bn != i1m && bn != i2m && bn != i3m) { m12[sQ[qp]] = -1; ++badOri; }  // :123-141
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

// Which matcher runs (same results, bit for bit; P.optMatchPath forces one for the tests):
//   1  k_match_cta   windows with a few dozen candidates per query (the demo's 100 px): all candidates of a chunk of queries
//                    in shared memory;
//   3  brute force   (match_bf.cu) windows that cover most of the image, or keypoint capacities beyond k_match_cta's
//                    shared memory: distance matrix sharded by query rows over all SMs, ordered resolve from top-k lists;
//   2  k_match       one warp per pair, the plain ordered walk (needs no scratch).
int launch_match(const Params& P, const MatchArgs& A, int nPairs, int pairStep, float nnratio, int checkOri,
                 int window, cudaStream_t st) {
  if (nPairs <= 0) return 0;
  const MatchSmem LY = match_smem_layout(A.stride);
  const bool ctaFits = A.stride < 65535 && LY.total <= 200 * 1024;
  const bool bfFits = A.bf.topk != nullptr && A.stride < 45000;
  // expected candidates per query: the level-0 keys (about a fifth of the capacity at scale 1.2) in a (2 window)^2 box
  const double frac = std::min(1.0, 4.0 * (double)window * window / ((double)std::max(1, P.bounds[1] - P.bounds[0]) * std::max(1, P.bounds[3] - P.bounds[2])));
  const bool wide = 0.22 * A.stride * frac > 256.0;
  int path = P.optMatchSerial ? 2 : P.optMatchPath;
  if (path == 0) path = (bfFits && (wide || !ctaFits)) ? 3 : ctaFits ? 1 : 2;
  if (path == 3 && !bfFits) path = ctaFits ? 1 : 2;
  if (path == 1 && !ctaFits) path = 2;
  if (path == 3) return launch_match_bf(P, A, A.bf, nPairs, pairStep, nnratio, checkOri, window, st);
  if (path == 1) {
    if (cudaFuncSetAttribute(k_match_cta, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) return 0;
    k_match_cta<<<nPairs, MT, LY.total, st>>>(P, A, pairStep, nnratio, checkOri, window);
  } else {
    // keypoint capacities beyond what the shared-memory tables hold: the one-warp-per-pair walk
    k_match<<<nPairs, 32, 0, st>>>(P, A, pairStep, nnratio, checkOri, window);
  }
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
// The same query for a BATCH of windows on frames that live on the device (what the tracking stage after initialisation
// asks for, tracking.cpp:44-47: radius searches around projected points with a level range): one warp per query,
// results in the reference's visiting order.  q = (x, y, r, minLevel, maxLevel) per query, slot = frame of each query.
__global__ void __launch_bounds__(256) k_features_in_area_batch(const __grid_constant__ Params P, const orbx_keypoint* __restrict__ keysUn,
                                                                const u32* __restrict__ gridOff, const u32* __restrict__ gridIdx,
                                                                int stride, const int* __restrict__ slots, const float* __restrict__ q5,
                                                                int nQ, int* out, int cap, int* counts) {
  const int w = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (w >= nQ) return;
  const int slot = slots[w];
  const float x = q5[5 * w], y = q5[5 * w + 1], r = q5[5 * w + 2];
  const int minLevel = (int)q5[5 * w + 3], maxLevel = (int)q5[5 * w + 4];
  const orbx_keypoint* __restrict__ kun = keysUn + (size_t)slot * stride;
  const u32* __restrict__ gOff = gridOff + (size_t)slot * (kGridCells + 1);
  const u32* __restrict__ gIdx = gridIdx + (size_t)slot * stride;
  int* o = out + (size_t)w * cap;
  int n = 0, minCX, maxCX, minCY, maxCY;
  if (cell_rect(P, x, y, r, minCX, maxCX, minCY, maxCY)) {
    const bool check = (minLevel > 0) || (maxLevel >= 0);  // Frame.cpp:187
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
          if (p < cap) o[p] = i2;
        }
        n += __popc(m);
      }
    }
  }
  if (lane == 0) counts[w] = n;
}
void launch_features_in_area_batch(const Params& P, const orbx_keypoint* keysUn, const u32* gridOff, const u32* gridIdx, int stride,
                                   const int* slots, const float* q5, int nQ, int* out, int cap, int* counts, cudaStream_t st) {
  if (nQ > 0) k_features_in_area_batch<<<(nQ + 7) / 8, 256, 0, st>>>(P, keysUn, gridOff, gridIdx, stride, slots, q5, nQ, out, cap, counts);
}

void launch_features_in_area(const Params& P, const orbx_keypoint* kun, const u32* gOff, const u32* gIdx, float x,
                             float y, float r, int minLevel, int maxLevel, int* out, int cap, int* count,
                             cudaStream_t st) {
  k_features_in_area<<<1, 32, 0, st>>>(P, kun, gOff, gIdx, x, y, r, minLevel, maxLevel, out, cap, count);
}

}  // namespace orbx
