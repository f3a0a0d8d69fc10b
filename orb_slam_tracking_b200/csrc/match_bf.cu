// ORBmatcher::SearchForInitialization (Features/ORBmatcher.cpp:11-150) for LARGE candidate sets: windows that cover most
// of the image (BASELINE.json's brute-force configuration: every level-0 key of F2 is a candidate of every level-0 key of
// F1) and keypoint capacities whose tables do not fit k_match_cta's shared memory.  Same results as the reference, bit for
// bit; what changes is where the Hamming distances are computed.
//
// The reference loop over i1 is order-dependent only through `vMatchedDistance[i2] <= dist` (:67): a candidate that an
// EARLIER query matched at least as well is skipped.  Which candidates a query sees, in which order, and at which distance
// is independent of the other queries.  So:
//   k_bf_prep     per pair: the level-0 keys of F2 in the grid's visiting order (cell ix, then iy, then insertion order --
//                 the order Frame::GetFeaturesInArea returns them in, SlamTypes/Frame.cpp:163-206) and the level-0 keys of
//                 F1 in ascending i1 (:36-39).
//   k_bf_topk     the distance matrix, sharded by QUERY ROWS over CTAs (and over GPUs: rows [q0, q1) per rank): a warp per
//                 query, lanes over the candidates, two 128-bit loads and 8 x POPC per pair; each lane keeps its three
//                 smallest (distance, visiting position) keys and the warp merges them into the query's sorted list of the
//                 kBfK smallest -- exact as long as no lane contributes three of them, which the list length records.
//   k_bf_resolve  per pair, one warp walks the queries in order and replays the reference's bookkeeping from the lists:
//                 the first two entries that are not skipped are bestDist / bestDist2 (:70-76).  Only a query whose list
//                 runs out before two such entries are found (its best candidates were taken by earlier queries) is
//                 rescanned against all candidates, with the skip rule applied, exactly like k_match does.
// The same top-k kernel without geometry is the raw all-pairs best / second-best benchmark (orbx_hamming_best2).
#include <limits.h>

#include "match_common.cuh"

namespace orbx {

constexpr int kBfThreads = 256, kBfWarps = kBfThreads / 32;
constexpr int kBfQPerCta = 16;          // queries per CTA of k_bf_topk (2 per warp): 500 CTAs for 8000 queries, small tail
constexpr u32 kNoKey = 0xffffffffu;

__global__ void __launch_bounds__(kBfThreads) k_bf_prep(const __grid_constant__ Params P, const MatchArgs A, const BfScratch B, int pairStep) {
  __shared__ int sPart[kBfThreads];
  const int pair = blockIdx.x, tid = threadIdx.x;
  const int s1 = pair * pairStep, s2 = s1 + 1;
  const int n1 = A.n[s1];
  const orbx_keypoint* __restrict__ k1 = A.keysUn + (size_t)s1 * A.stride;
  const orbx_keypoint* __restrict__ k2 = A.keysUn + (size_t)s2 * A.stride;
  const u32* __restrict__ gOff = A.gridOff + (size_t)s2 * (kGridCells + 1);
  const u32* __restrict__ gIdx = A.gridIdx + (size_t)s2 * A.stride;
  u32* cand = B.cand + (size_t)pair * A.stride;
  float2* candXY = B.candXY + (size_t)pair * A.stride;
  int* qIdx = B.qIdx + (size_t)pair * A.stride;
  // level-0 keys of F2 in visiting order (thread t owns cells [12 t, 12 t + 12))
  constexpr int CPT = kGridCells / kBfThreads;
  static_assert(CPT * kBfThreads == kGridCells, "cells per thread");
  const int c0 = tid * CPT;
  int tot = 0;
  for (u32 e = gOff[c0]; e < gOff[c0 + CPT]; ++e) tot += (k2[gIdx[e]].octave == 0);
  sPart[tid] = tot;
  __syncthreads();
  for (int o = 1; o < kBfThreads; o <<= 1) {
    const int v = tid >= o ? sPart[tid - o] : 0;
    __syncthreads();
    sPart[tid] += v;
    __syncthreads();
  }
  {
    int run = tid ? sPart[tid - 1] : 0;
    u32 e = gOff[c0];
    for (int j = 0; j < CPT; ++j) {
      const u32 e1 = gOff[c0 + j + 1];
      const u32 cell = (u32)(c0 + j), ix = cell / kGridRows, iy = cell - ix * kGridRows;
      for (; e < e1; ++e) {
        const int i2 = (int)gIdx[e];
        const orbx_keypoint k = k2[i2];
        if (k.octave == 0) { cand[run] = (u32)i2 | (ix << 16) | (iy << 22); candXY[run] = make_float2(k.x, k.y); ++run; }
      }
    }
    if (tid == kBfThreads - 1) B.counts[pair * 4 + 0] = run;
  }
  __syncthreads();
  // level-0 queries of F1, ascending i1 (:36-39)
  {
    const int per = (n1 + kBfThreads - 1) / kBfThreads, b = min(tid * per, n1), e = min(b + per, n1);
    int c = 0;
    for (int i = b; i < e; ++i) c += !(k1[i].octave > 0);
    sPart[tid] = c;
    __syncthreads();
    for (int o = 1; o < kBfThreads; o <<= 1) {
      const int v = tid >= o ? sPart[tid - o] : 0;
      __syncthreads();
      sPart[tid] += v;
      __syncthreads();
    }
    int run = tid ? sPart[tid - 1] : 0;
    for (int i = b; i < e; ++i)
      if (!(k1[i].octave > 0)) qIdx[run++] = i;
    if (tid == kBfThreads - 1) B.counts[pair * 4 + 1] = run;
  }
}

// the three smallest keys of a lane, ascending; branch-free insertion
struct Top3 {
  u32 k0, k1, k2;
  __device__ __forceinline__ void init() { k0 = k1 = k2 = kNoKey; }
  __device__ __forceinline__ void insert(u32 k) {
    const u32 n2 = min(k2, max(k1, k)), n1 = min(k1, max(k0, k));
    k0 = min(k0, k); k1 = n1; k2 = n2;
  }
};

// Merges the lanes' Top3 into the warp's sorted list of the kBfK smallest keys.  Returns the number of entries that are
// exact: the merge stops when a lane has given all three of its keys (its fourth is unknown), or when the keys run out.
__device__ __forceinline__ int merge_top(Top3& T, u32* out, int lane) {
  int given = 0, n = 0;
#pragma unroll
  for (int j = 0; j < kBfK; ++j) {
    const u32 best = __reduce_min_sync(0xffffffffu, T.k0);
    if (best == kNoKey) break;
    if (lane == 0) out[n] = best;
    ++n;
    bool dry = false;
    if (T.k0 == best) {  // keys are unique (they carry the visiting position)
      T.k0 = T.k1; T.k1 = T.k2; T.k2 = kNoKey;
      dry = ++given == 3;
    }
    if (__any_sync(0xffffffffu, dry)) break;
  }
  return n;
}

// Row block [q0, q1) of the distance matrix of every pair: per query the sorted list of its kBfK best candidates.
// topk row of a query: kBfK keys ((distance << 20) | visiting position), then one word:
//   bits 0..19 number of candidates of the query (GetFeaturesInArea's result size), bits 24..27 number of exact list entries.
__global__ void __launch_bounds__(kBfThreads) k_bf_topk(const __grid_constant__ Params P, const MatchArgs A, const BfScratch B, int pairStep,
                                                       int window, int q0, int q1) {
  const int pair = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nC = B.counts[pair * 4 + 0], nQ = min(B.counts[pair * 4 + 1], q1);
  const int qBase = q0 + blockIdx.x * kBfQPerCta;
  if (qBase >= nQ) return;
  const int s1 = pair * pairStep, s2 = s1 + 1;
  const orbx_keypoint* __restrict__ k1 = A.keysUn + (size_t)s1 * A.stride;
  const u8* __restrict__ d1 = A.desc + (size_t)s1 * A.stride * 32;
  const u8* __restrict__ d2 = A.desc + (size_t)s2 * A.stride * 32;
  const u32* __restrict__ cand = B.cand + (size_t)pair * A.stride;
  const float2* __restrict__ candXY = B.candXY + (size_t)pair * A.stride;
  const int* __restrict__ qIdx = B.qIdx + (size_t)pair * A.stride;
  u32* topk = B.topk + (size_t)pair * A.stride * (kBfK + 1);
  const float r = (float)window;
  for (int q = qBase + warp; q < min(qBase + kBfQPerCta, nQ); q += kBfWarps) {
    const int i1 = qIdx[q];
    const orbx_keypoint kp1 = k1[i1];
    const float x = kp1.x, y = kp1.y;
    int minCX, maxCX, minCY, maxCY;
    const bool any = cell_rect(P, x, y, r, minCX, maxCX, minCY, maxCY);
    Top3 T;
    T.init();
    int cnt = 0;
    if (any) {
      const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32));
      const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32) + 1);
      for (int t = lane; t < nC; t += 32) {
        const u32 c = __ldg(cand + t);
        const int ix = (int)((c >> 16) & 63), iy = (int)(c >> 22);
        if (ix < minCX || ix > maxCX || iy < minCY || iy > maxCY) continue;  // cells GetFeaturesInArea visits, Frame.cpp:167-190
        const float2 p = __ldg(candXY + t);
        if (!(fabsf(p.x - x) < r && fabsf(p.y - y) < r)) continue;            // Frame.cpp:196-199
        const int i2 = (int)(c & 0xffffu);
        ++cnt;
        T.insert(((u32)hamming256(a0, a1, d2 + (size_t)i2 * 32) << 20) | (u32)t);
      }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    u32* row = topk + (size_t)q * (kBfK + 1);
    const int n = merge_top(T, row, lane);
    if (lane == 0) {
      for (int j = n; j < kBfK; ++j) row[j] = kNoKey;
      row[kBfK] = (u32)cnt | ((u32)n << 24);
    }
  }
}

// One warp per pair: the reference's loop over the queries, in order (:36-117), from the lists of k_bf_topk.
// Batches of 32 queries: first every lane stages one query's list (its keys translated to keypoint indices, the
// candidates' angles beside them) into shared memory with independent loads; then the queries are resolved one after
// the other from shared memory only -- lane j looks at list entry j, one ballot finds the first two entries that :67
// does not skip.  The serial chain never waits for global memory except in the rare full rescan.
__global__ void __launch_bounds__(32) k_bf_resolve(const __grid_constant__ Params P, const MatchArgs A, const BfScratch B, int pairStep,
                                                   float nnratio, int checkOri, int window) {
  extern __shared__ __align__(16) u8 dsm[];
  __shared__ int sHist[HISTO_LENGTH];
  __shared__ u32 sEnt[32][8];     // per query of the batch: 7 x (distance << 16 | i2), info word
  __shared__ float sAng[32][8];   // angle of each listed candidate, [7] = the query's own angle
  __shared__ int sI1[32];
  const int pair = blockIdx.x, lane = threadIdx.x;
  const int s1 = pair * pairStep, s2 = s1 + 1;
  const int n1 = A.n[s1], n2 = A.n[s2];
  const int nC = B.counts[pair * 4 + 0], nQ = B.counts[pair * 4 + 1];
  unsigned short* const sMD = reinterpret_cast<unsigned short*>(dsm);   // vMatchedDistance (0xffff = INT_MAX; distances are <= 256)
  unsigned short* const sM21 = sMD + A.stride;                           // vnMatches21 (0xffff = -1)
  u32* const sTouch = reinterpret_cast<u32*>(sM21 + A.stride);           // per key of F2: the first query (lane) of the running batch that takes it
  u8* const sBin = reinterpret_cast<u8*>(sTouch + A.stride);             // rotation bin of query position q (0xff = none)
  const orbx_keypoint* __restrict__ k1 = A.keysUn + (size_t)s1 * A.stride;
  const orbx_keypoint* __restrict__ k2 = A.keysUn + (size_t)s2 * A.stride;
  const u8* __restrict__ d1 = A.desc + (size_t)s1 * A.stride * 32;
  const u8* __restrict__ d2 = A.desc + (size_t)s2 * A.stride * 32;
  const u32* __restrict__ cand = B.cand + (size_t)pair * A.stride;
  const float2* __restrict__ candXY = B.candXY + (size_t)pair * A.stride;
  const int* __restrict__ qIdx = B.qIdx + (size_t)pair * A.stride;
  const u32* __restrict__ topk = B.topk + (size_t)pair * A.stride * (kBfK + 1);
  int* m12 = A.matches12 + (size_t)pair * A.stride;
  for (int i = lane; i < n1; i += 32) m12[i] = -1;
  for (int i = lane; i < n2; i += 32) { sMD[i] = 0xffffu; sM21[i] = 0xffffu; sTouch[i] = 0xffffffffu; }
  for (int i = lane; i < nQ; i += 32) sBin[i] = 0xff;
  if (lane < HISTO_LENGTH) sHist[lane] = 0;
  __syncwarp();
  const float r = (float)window;
  const float factor = (float)HISTO_LENGTH / 360.0f;
  int nmatches = 0, badDist = 0, badRatio = 0;

  int qb = 0;
  // the reference's loop body for query qi of the staged batch against the current state, lanes over its list (and, when
  // the list runs out, over all candidates); returns the matched key of F2 or -1
  auto resolve_one = [&](int qi) -> int {
      const u32 info = sEnt[qi][7];
      const int nCand = (int)(info & 0xfffffu), nList = (int)(info >> 24);
      if (nCand == 0) return -1;  // vIndices2.empty(), :46
      // lane j < nList: is list entry j skipped by :67 ?
      u32 e = kNoKey;
      bool ok = false;
      if (lane < nList) {
        e = sEnt[qi][lane];
        ok = !((int)sMD[e & 0xffffu] <= (int)(e >> 16));
      }
      const u32 live = __ballot_sync(0xffffffffu, ok);
      const int j0 = __ffs((int)live) - 1, j1 = __ffs((int)(live & (live - 1))) - 1;
      int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx = -1;
      float ang2 = 0.f;
      if (j0 >= 0) {
        const u32 e0 = __shfl_sync(0xffffffffu, e, j0);
        bestDist = (int)(e0 >> 16); bestIdx = (int)(e0 & 0xffffu);
        ang2 = sAng[qi][j0];
      }
      if (j1 >= 0) bestDist2 = (int)(__shfl_sync(0xffffffffu, e, j1) >> 16);
      const int i1 = sI1[qi];
      if (j1 < 0 && nList < nCand) {
        // the list ran out before the candidates did: the ordered walk over all candidates of the query
        const orbx_keypoint kp1 = k1[i1];
        int minCX, maxCX, minCY, maxCY;
        cell_rect(P, kp1.x, kp1.y, r, minCX, maxCX, minCY, maxCY);
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32));
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(d1 + (size_t)i1 * 32) + 1);
        u32 key = kNoKey;
        int bidx = -1, bd2 = INT_MAX;
        for (int t = lane; t < nC; t += 32) {
          const u32 c = __ldg(cand + t);
          const int ix = (int)((c >> 16) & 63), iy = (int)(c >> 22);
          if (ix < minCX || ix > maxCX || iy < minCY || iy > maxCY) continue;
          const float2 p = __ldg(candXY + t);
          if (!(fabsf(p.x - kp1.x) < r && fabsf(p.y - kp1.y) < r)) continue;
          const int i2 = (int)(c & 0xffffu);
          const int dist = hamming256(a0, a1, d2 + (size_t)i2 * 32);
          if ((int)sMD[i2] <= dist) continue;  // :67
          const u32 k = ((u32)dist << 20) | (u32)t;
          if (k < key) {
            if (key != kNoKey) bd2 = (int)(key >> 20);
            key = k; bidx = i2;
          } else if (dist < bd2) bd2 = dist;
        }
        const u32 best = __reduce_min_sync(0xffffffffu, key);
        const int winner = __ffs((int)__ballot_sync(0xffffffffu, key == best)) - 1;
        bestDist = best == kNoKey ? INT_MAX : (int)(best >> 20);
        const int mine2 = (lane == winner) ? bd2 : (key == kNoKey ? INT_MAX : (int)(key >> 20));
        bestDist2 = (int)__reduce_min_sync(0xffffffffu, (u32)mine2);
        bestIdx = __shfl_sync(0xffffffffu, bidx, winner);
        if (bestIdx >= 0) ang2 = k2[bestIdx].angle;
      }
      if (bestDist > TH_LOW) { ++badDist; return -1; }                                        // :82
      if ((float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) { ++badRatio; return -1; }  // :87
      const int prev = sM21[bestIdx] == 0xffffu ? -1 : (int)sM21[bestIdx];
      if (prev >= 0) --nmatches;  // :95-98
      ++nmatches;
      __syncwarp();
      if (lane == 0) {
        if (prev >= 0) m12[prev] = -1;
        m12[i1] = bestIdx;
        sM21[bestIdx] = (unsigned short)i1;
        sMD[bestIdx] = (unsigned short)bestDist;
        if (checkOri) {
          float rot = sAng[qi][7] - ang2;
          if (rot < 0.0f) rot = rot + 360.0f;
          int bin = (int)roundf(__fmul_rn(rot, factor));
          if (bin == HISTO_LENGTH) bin = 0;
          sBin[qb + qi] = (u8)bin;
          sHist[bin] += 1;
        }
      }
      __syncwarp();
      return bestIdx;
  };

  for (qb = 0; qb < nQ; qb += 32) {
    {  // stage the batch
      const int q = qb + lane;
      u32 w[8] = {kNoKey, kNoKey, kNoKey, kNoKey, kNoKey, kNoKey, kNoKey, 0u};
      int i1 = 0;
      if (q < nQ) {
        const uint4 lo = __ldg(reinterpret_cast<const uint4*>(topk + (size_t)q * 8)), hi = __ldg(reinterpret_cast<const uint4*>(topk + (size_t)q * 8) + 1);
        w[0] = lo.x; w[1] = lo.y; w[2] = lo.z; w[3] = lo.w; w[4] = hi.x; w[5] = hi.y; w[6] = hi.z; w[7] = hi.w;
        i1 = __ldg(qIdx + q);
      }
      const int nList = (int)(w[7] >> 24);
#pragma unroll
      for (int j = 0; j < kBfK; ++j) {
        u32 e = kNoKey;
        float ang = 0.f;
        if (j < nList) {
          const int i2 = (int)(__ldg(cand + (w[j] & 0xfffffu)) & 0xffffu);
          e = ((w[j] >> 20) << 16) | (u32)i2;
          ang = k2[i2].angle;
        }
        sEnt[lane][j] = e;
        sAng[lane][j] = ang;
      }
      sEnt[lane][7] = w[7];
      sAng[lane][7] = q < nQ ? k1[i1].angle : 0.f;
      sI1[lane] = i1;
    }
    __syncwarp();
    // Every lane decides ITS query of the batch from its list against the state at the start of the batch; a lane whose best
    // or second-best key is taken by an earlier lane of the batch, or whose list runs out, is decided again by the whole warp
    // (resolve_one) once every lane before it has committed.  See k_match_cta, phase (d): same argument, same result as the
    // reference's ordered loop.
    {
      const u32 info = sEnt[lane][7];
      const int nCand = (int)(info & 0xfffffu), nList = (int)(info >> 24);
      int best = INT_MAX, best2 = INT_MAX, bidx = -1, bidx2 = -1, j0 = 0;
#pragma unroll
      for (int j = 0; j < kBfK; ++j) {
        if (j < nList) {
          const u32 e = sEnt[lane][j];
          const int dist = (int)(e >> 16), i2 = (int)(e & 0xffffu);
          if (!((int)sMD[i2] <= dist)) {  // :67; the list is sorted (distance, visiting position)
            if (bidx < 0) { best = dist; bidx = i2; j0 = j; }
            else if (bidx2 < 0) { best2 = dist; bidx2 = i2; }
          }
        }
      }
      const bool hard = bidx2 < 0 && nList < nCand;  // the list ran out before the candidates did
      const bool bad = best > TH_LOW;
      const bool ratio = !bad && (float)best > __fmul_rn(nnratio, (float)best2);
      u32 pend = __ballot_sync(0xffffffffu, qb + lane < nQ && nCand > 0);
      int myTouch = -1;
      while (pend) {
        const bool mine = (pend >> lane) & 1u;
        const bool isMatch = mine && !hard && !bad && !ratio;
        if (isMatch) atomicMin(&sTouch[bidx], (u32)lane);
        __syncwarp();
        const bool conf = mine && (hard || (!bad && (sTouch[bidx] < (u32)lane || (bidx2 >= 0 && sTouch[bidx2] < (u32)lane))));
        const u32 cm = __ballot_sync(0xffffffffu, conf);
        const int f = cm ? __ffs((int)cm) - 1 : 32;
        const u32 commit = f >= 32 ? pend : pend & ((1u << f) - 1u);
        const bool doCommit = (commit >> lane) & 1u;
        __syncwarp();
        if (isMatch && !doCommit && sTouch[bidx] == (u32)lane) sTouch[bidx] = 0xffffffffu;  // tentative claims are withdrawn
        int prev = -1;
        if (doCommit && isMatch) {
          prev = sM21[bidx] == 0xffffu ? -1 : (int)sM21[bidx];
          const int i1 = sI1[lane];
          if (prev >= 0) m12[prev] = -1;  // :95-98
          m12[i1] = bidx;
          sM21[bidx] = (unsigned short)i1;
          sMD[bidx] = (unsigned short)best;
          myTouch = bidx;
          if (checkOri) {
            float rot = sAng[lane][7] - sAng[lane][j0];
            if (rot < 0.0f) rot = rot + 360.0f;
            int bin = (int)roundf(__fmul_rn(rot, factor));
            if (bin == HISTO_LENGTH) bin = 0;
            sBin[qb + lane] = (u8)bin;
            atomicAdd(&sHist[bin], 1);
          }
        }
        badDist += __popc(__ballot_sync(0xffffffffu, doCommit && bad));
        badRatio += __popc(__ballot_sync(0xffffffffu, doCommit && ratio));
        nmatches += __popc(__ballot_sync(0xffffffffu, doCommit && isMatch)) - __popc(__ballot_sync(0xffffffffu, prev >= 0));
        pend &= ~commit;
        __syncwarp();
        if (f < 32) {
          const int got = resolve_one(f);
          if (got >= 0) {
            if (lane == 0) atomicMin(&sTouch[got], (u32)f);
            if (lane == f) myTouch = got;
          }
          pend &= ~(1u << f);
          __syncwarp();
        }
      }
      if (myTouch >= 0) sTouch[myTouch] = 0xffffffffu;
    }
    __syncwarp();
  }

  int badOri = 0;
  if (checkOri) {
    int i1m, i2m, i3m;
    three_maxima(sHist, i1m, i2m, i3m);  // ComputeThreeMaxima, :152-183
    for (int qp = lane; qp < nQ; qp += 32) {
      const int bn = sBin[qp];
      if (bn != 0xff && bn != i1m && bn != i2m && bn != i3m) { m12[qIdx[qp]] = -1; ++badOri; }  // :123-141
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

size_t bf_resolve_smem(int stride) { return (size_t)stride * 9 + 16; }

int launch_match_bf(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, float nnratio, int checkOri,
                    int window, cudaStream_t st) {
  k_bf_prep<<<nPairs, kBfThreads, 0, st>>>(P, A, B, pairStep);
  dim3 grid((A.stride + kBfQPerCta - 1) / kBfQPerCta, nPairs);
  k_bf_topk<<<grid, kBfThreads, 0, st>>>(P, A, B, pairStep, window, 0, INT_MAX);
  const size_t smem = bf_resolve_smem(A.stride);
  if (smem > 48 * 1024 && cudaFuncSetAttribute(k_bf_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
  k_bf_resolve<<<nPairs, 32, smem, st>>>(P, A, B, pairStep, nnratio, checkOri, window);
  return 3;
}

// the three steps separately: a rank computes the rows [q0, q1) of the distance matrix of a pair, any rank resolves
int launch_bf_prep(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, cudaStream_t st) {
  k_bf_prep<<<nPairs, kBfThreads, 0, st>>>(P, A, B, pairStep);
  return 1;
}
int launch_bf_rows(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, int window, int q0, int q1,
                   cudaStream_t st) {
  if (q1 <= q0) return 0;
  dim3 grid((std::min(q1, A.stride) - q0 + kBfQPerCta - 1) / kBfQPerCta, nPairs);
  k_bf_topk<<<grid, kBfThreads, 0, st>>>(P, A, B, pairStep, window, q0, q1);
  return 1;
}
int launch_bf_resolve(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, float nnratio, int checkOri,
                      int window, cudaStream_t st) {
  const size_t smem = bf_resolve_smem(A.stride);
  if (smem > 48 * 1024 && cudaFuncSetAttribute(k_bf_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
  k_bf_resolve<<<nPairs, 32, smem, st>>>(P, A, B, pairStep, nnratio, checkOri, window);
  return 1;
}

// Raw all-pairs Hamming (the micro-benchmark of BASELINE.json's brute-force configuration): for each of nA descriptors
// the sorted list of its kBfK nearest of nB (ties by index, like the matcher's strict `<`), no geometry.  Rows as above.
__global__ void __launch_bounds__(kBfThreads) k_hamming_topk(const u8* __restrict__ a, int nA, const u8* __restrict__ b, int nB, u32* topk) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int qBase = blockIdx.x * kBfQPerCta;
  for (int q = qBase + warp; q < min(qBase + kBfQPerCta, nA); q += kBfWarps) {
    const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(a + (size_t)q * 32));
    const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(a + (size_t)q * 32) + 1);
    Top3 T;
    T.init();
    for (int t = lane; t < nB; t += 32) T.insert(((u32)hamming256(a0, a1, b + (size_t)t * 32) << 20) | (u32)t);
    u32* row = topk + (size_t)q * (kBfK + 1);
    const int n = merge_top(T, row, lane);
    if (lane == 0) {
      for (int j = n; j < kBfK; ++j) row[j] = kNoKey;
      row[kBfK] = (u32)nB | ((u32)n << 24);
    }
  }
}
int launch_hamming_topk(const u8* a, int nA, const u8* b, int nB, u32* topk, cudaStream_t st) {
  if (nA <= 0) return 0;
  k_hamming_topk<<<(nA + kBfQPerCta - 1) / kBfQPerCta, kBfThreads, 0, st>>>(a, nA, b, nB, topk);
  return 1;
}

}  // namespace orbx
