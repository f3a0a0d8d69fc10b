// ComputeKeyPointsOld (Features/ORBextractor.cpp:1207-1496): the selection policy the reference keeps
// beside ComputeKeyPointsOctTree (its call is commented out at :1555).  k_fast produced the FAST keypoints of the policy's own
// cell grid (minThFAST retry for cells with at most 3 keypoints); this kernel does the rest for one (level, slot) per CTA:
//   1  the candidates into the reference's order -- cells row-major, inside a cell the order cv::FAST emits (y, then x).
//      Their emission-order keys are unique, so a keypoint's position is the number of keys below its own: counted
//      against shared-memory chunks of all keys (no sorting network, any candidate count);
//   2  per cell nTotal, then the redistribution of the per-cell quotas (:1376-1425), a few hundred integers, one thread;
//   3  cv::KeyPointsFilter::retainBest per cell (:1452): the first nToRetain elements after libstdc++'s
//      std::nth_element (stdnth_emul_t, checked on the host against the real one) -- one thread per cell;
//   4  concatenation in cell order and, if the level still holds more than its quota, retainBest over the level (:1480-1484).
// Output: the level's keypoint slots (lvlXY / lvlResp / lvlCount), which orientation, blur and descriptors consume unchanged.
#include "common.cuh"

namespace orbx {

constexpr int kOldThreads = 256, kOldChunk = 2048;

struct RespGreater {  // retainBest's comparator on the packed record: response in bits 48..63
  __host__ __device__ __forceinline__ bool operator()(const u64& a, const u64& b) const { return (a >> 48) > (b >> 48); }
};

__global__ void __launch_bounds__(kOldThreads) k_retain_old(const __grid_constant__ Params P) {
  __shared__ u32 sOrd[kOldChunk];
  extern __shared__ int sCell[];  // [nCells] count | [nCells + 1] start | [nCells] keep
  __shared__ int sTotal;
  const int l = blockIdx.x, slot = blockIdx.y, tid = threadIdx.x;
  const LevelInfo& L = P.lv[l];
  const int nCells = L.oCols * L.oRows;
  int* const cCnt = sCell;
  int* const cStart = sCell + nCells;
  int* const cKeep = cStart + nCells + 1;
  const size_t co = (size_t)slot * P.candTotal + L.candOff;
  const int M = min((int)P.candCount[(size_t)slot * kMaxLevels + l], L.candCap);
  u64* const key = P.candKey + co;          // (response - ... ) << 32 | ~order; reused below as the level's gather buffer
  const u32* const xy = P.candXY + co;
  u32* const rank = P.candNode + co;        // free under this policy: the position of each candidate
  u64* const sorted = P.oldSorted + co;
  const int nDesired = L.quota;
  int* const lvlCount = P.lvlCount + (size_t)slot * kMaxLevels + l;
  const size_t ko = (size_t)slot * P.kpStride + L.kpOff;

  for (int c = tid; c < nCells; c += kOldThreads) cCnt[c] = 0;
  for (int i = tid; i < M; i += kOldThreads) rank[i] = 0;
  __syncthreads();
  // 1: cell populations and positions
  for (int i = tid; i < M; i += kOldThreads) atomicAdd(&cCnt[(~(u32)key[i]) >> 18], 1);
  for (int c0 = 0; c0 < M; c0 += kOldChunk) {
    const int cn = min(kOldChunk, M - c0);
    __syncthreads();
    for (int j = tid; j < cn; j += kOldThreads) sOrd[j] = ~(u32)key[c0 + j];
    __syncthreads();
    for (int i = tid; i < M; i += kOldThreads) {
      const u32 o = ~(u32)key[i];
      int r = 0;
      for (int j = 0; j < cn; ++j) r += sOrd[j] < o;
      rank[i] += (u32)r;
    }
  }
  __syncthreads();
  for (int i = tid; i < M; i += kOldThreads) {
    const u64 k = key[i];
    const u32 p = xy[i];
    // response + 1 was the segment-test value; cv::FAST's response is value - 1, already what the key holds
    sorted[rank[i]] = ((k >> 32) << 48) | (u64)p;
  }
  // 2: quotas per cell (one thread; a few hundred cells at most)
  if (tid == 0) {
    int run = 0;
    for (int c = 0; c < nCells; ++c) { cStart[c] = run; run += cCnt[c]; }
    cStart[nCells] = run;
    const int maxBX = L.w - kEdge, maxBY = L.h - kEdge;
    const bool skipRow = maxBY + 3 - (kEdge + (L.oRows - 1) * L.oCellH - 3) <= 0;   // hY <= 0, :1300
    const bool skipCol = maxBX + 3 - (kEdge + (L.oCols - 1) * L.oCellW - 3) <= 0;   // hX <= 0, :1319
    const int nf = L.oNfCell;
    int nNoMore = 0, nToDistribute = 0;
    // cKeep doubles as the bNoMore flag during the loop: bit 30
    for (int c = 0; c < nCells; ++c) {
      const int i = c / L.oCols, j = c - i * L.oCols;
      if ((skipRow && i == L.oRows - 1) || (skipCol && j == L.oCols - 1)) { cKeep[c] = 0; continue; }  // never visited: bNoMore stays false
      const int n = cCnt[c];
      if (n > nf) cKeep[c] = nf;
      else { cKeep[c] = n | (1 << 30); nToDistribute += nf - n; ++nNoMore; }
    }
    while (nToDistribute > 0 && nNoMore < nCells) {  // :1390-1425
      const int nNew = nf + (int)ceilf(__fdiv_rn((float)nToDistribute, (float)(nCells - nNoMore)));
      nToDistribute = 0;
      for (int c = 0; c < nCells; ++c) {
        if (cKeep[c] & (1 << 30)) continue;
        const int n = cCnt[c];
        if (n > nNew) cKeep[c] = nNew;
        else { cKeep[c] = n | (1 << 30); nToDistribute += nNew - n; ++nNoMore; }
      }
    }
    int total = 0;
    for (int c = 0; c < nCells; ++c) { cKeep[c] &= ~(1 << 30); total += cKeep[c]; }
    sTotal = total;
  }
  __syncthreads();
  // 3: retainBest per cell: a thread per cell replays std::nth_element on the cell's records
  for (int c = tid; c < nCells; c += kOldThreads) {
    const int n = cCnt[c], keep = cKeep[c];
    if (keep > 0 && n > keep) stdnth_emul_t(sorted + cStart[c], (long)n, (long)(keep - 1), RespGreater());
  }
  __syncthreads();
  // 4: the level's list = kept records in cell order
  const int total = sTotal;
  if (total > L.candCap) {  // cannot happen (kept <= candidates); guards the gather buffer
    if (tid == 0) { atomicMin(P.status + slot, ORBX_ERR_CAPACITY); *lvlCount = 0; }
    return;
  }
  __syncthreads();
  if (tid == 0) {  // prefix of the kept counts into cCnt (the populations are no longer needed)
    int run = 0;
    for (int c = 0; c < nCells; ++c) { const int k = cKeep[c]; cCnt[c] = run; run += k; }
  }
  __syncthreads();
  const bool cut = total > nDesired;
  for (int c = 0; c < nCells; ++c) {
    const int k = cKeep[c], dst = cCnt[c], src = cStart[c];
    for (int i = tid; i < k; i += kOldThreads) {
      const u64 r = sorted[src + i];
      if (cut) key[dst + i] = r;  // gather for the level-wide retainBest
      else {
        const u32 p = (u32)r;
        P.lvlXY[ko + dst + i] = ((p & 0xffffu) + kMinBorder) | (((p >> 16) + kMinBorder) << 16);
        P.lvlResp[ko + dst + i] = (float)(u32)(r >> 48);
      }
    }
  }
  __syncthreads();
  if (cut) {
    if (tid == 0) stdnth_emul_t(key, (long)total, (long)(nDesired - 1), RespGreater());  // :1480-1484
    __syncthreads();
    for (int i = tid; i < nDesired; i += kOldThreads) {
      const u64 r = key[i];
      const u32 p = (u32)r;
      P.lvlXY[ko + i] = ((p & 0xffffu) + kMinBorder) | (((p >> 16) + kMinBorder) << 16);
      P.lvlResp[ko + i] = (float)(u32)(r >> 48);
    }
  }
  if (tid == 0) *lvlCount = min(total, nDesired);
}

int launch_retain_old(const Params& P, int nSlots, cudaStream_t st) {
  int maxCells = 1;
  for (int l = 0; l < P.nlevels; ++l) maxCells = std::max(maxCells, P.lv[l].oCols * P.lv[l].oRows);
  dim3 grid(P.nlevels, nSlots);
  k_retain_old<<<grid, kOldThreads, (size_t)(3 * maxCells + 1) * sizeof(int), st>>>(P);
  return 1;
}

}  // namespace orbx
