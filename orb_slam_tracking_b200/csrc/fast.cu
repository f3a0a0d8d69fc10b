// Per-cell FAST-9/16 with 3x3 non-maximum suppression and the min-threshold retry
// (ORBextractor::ComputeKeyPointsOctTree cell loop, Features/ORBextractor.cpp:1051-1141, calling
// cv::FAST(roi, kps, th, true) at :1109 and :1119; arithmetic in SURVEY.md appendix A3).
//
// One CTA per (cell, slot); the cells of ALL levels are in one table, so one launch covers a batch.
// The segment-test value m(p) = max over the 16 arcs of 9 of min(v - ring) / min(ring - v) does not
// depend on the threshold: p is a corner at t iff m > t and its NMS score is m - 1.  A corner
// survives NMS at t iff m >= 2 and m is a strict 3x3 maximum of the m-map restricted to the
// cell's tested interior (a neighbour that is not a corner at t has m' <= t < m anyway).  So the
// CTA computes the peaks once and then applies the reference's rule: keep the peaks with
// m > iniTh, or, if the cell has none, the peaks with m > minTh.
//
// Survivors are appended to the level's candidate list with an atomic cursor; each carries its
// position in the reference's emission order (cell row-major, then y, then x) inside its key, which
// is all the quadtree needs ("first strict maximum" == max response, then min order).
#include "common.cuh"

namespace orbx {

constexpr int kStride = kMaxCellDim;

__device__ __forceinline__ int arc_value(const int (&d)[16]) {
  // max over 16 arcs of min over 9 consecutive ring differences, and the same for -d
  int mn2[16], mx2[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) { mn2[k] = min(d[k], d[(k + 1) & 15]); mx2[k] = max(d[k], d[(k + 1) & 15]); }
  int mn4[16], mx4[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) { mn4[k] = min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = max(mx2[k], mx2[(k + 2) & 15]); }
  int A = -256, B = -256;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int mn = min(min(mn4[k], mn4[(k + 4) & 15]), d[(k + 8) & 15]);
    const int mx = max(max(mx4[k], mx4[(k + 4) & 15]), d[(k + 8) & 15]);
    A = max(A, mn);
    B = max(B, -mx);
  }
  return max(A, B);
}

__global__ void __launch_bounds__(256) k_fast(const __grid_constant__ Params P) {
  __shared__ u8 sImg[kStride * kStride];
  __shared__ u8 sM[kStride * kStride];          // m-map of the interior, 1-px zero frame
  __shared__ unsigned short sList[kStride * kStride];
  __shared__ int sN, sPeaks, sIni, sMin, sBase, sCursor;

  const CellDesc c = P.cells[blockIdx.x];
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[c.level];
  const int cw = c.x1 - c.x0, ch = c.y1 - c.y0;
  if (cw < 7 || ch < 7) return;  // cv::FAST returns nothing for such ROIs
  const int iw = cw - 6, ih = ch - 6;
  const int tid = threadIdx.x;
  const u8* __restrict__ img = level_ptr((const u8*)P.pyr, P, slot, c.level) + (size_t)c.y0 * L.pitch + c.x0;

  for (int i = tid; i < cw * ch; i += 256) {
    const int r = i / cw, x = i - r * cw;
    sImg[r * kStride + x] = __ldg(img + (size_t)r * L.pitch + x);
  }
  for (int i = tid; i < (ih + 2) * kStride; i += 256) sM[i] = 0;
  if (tid == 0) { sN = 0; sPeaks = 0; sIni = 0; sMin = 0; sCursor = 0; }
  __syncthreads();

  const int tIni = min(max(P.iniTh, 0), 255), tMin = min(max(P.minTh, 0), 255);
  const int tq = min(tIni, tMin);
  // ring offsets in shared-memory units, cv::FAST order
  constexpr int o0 = 3 * kStride, o1 = 3 * kStride + 1, o2 = 2 * kStride + 2, o3 = kStride + 3, o4 = 3,
                o5 = -kStride + 3, o6 = -2 * kStride + 2, o7 = -3 * kStride + 1, o8 = -3 * kStride,
                o9 = -3 * kStride - 1, o10 = -2 * kStride - 2, o11 = -kStride - 3, o12 = -3,
                o13 = kStride - 3, o14 = 2 * kStride - 2, o15 = 3 * kStride - 1;

  // pass 1: opposing-pair rejection (every arc of 9 contains one pixel of each opposing pair)
  for (int i = tid; i < iw * ih; i += 256) {
    const int yy = i / iw, xx = i - yy * iw;
    const u8* p = sImg + (yy + 3) * kStride + xx + 3;
    const int v = p[0];
    const int lo = v - tq, hi = v + tq;
    bool pass = true;
#define ORBX_PAIR(a, b)                                              \
  {                                                                  \
    const int pa = p[a], pb = p[b];                                  \
    if (pa >= lo && pa <= hi && pb >= lo && pb <= hi) pass = false;  \
  }
    ORBX_PAIR(o0, o8) ORBX_PAIR(o4, o12)
    if (pass) { ORBX_PAIR(o2, o10) ORBX_PAIR(o6, o14) }
    if (pass) { ORBX_PAIR(o1, o9) ORBX_PAIR(o3, o11) ORBX_PAIR(o5, o13) ORBX_PAIR(o7, o15) }
#undef ORBX_PAIR
    if (pass) sList[atomicAdd(&sN, 1)] = (unsigned short)((yy << 8) | xx);
  }
  __syncthreads();
  const int n = sN;
  // pass 2: exact segment-test value of the surviving pixels
  for (int i = tid; i < n; i += 256) {
    const int e = sList[i], yy = e >> 8, xx = e & 255;
    const u8* p = sImg + (yy + 3) * kStride + xx + 3;
    const int v = p[0];
    int d[16];
    d[0] = v - p[o0]; d[1] = v - p[o1]; d[2] = v - p[o2]; d[3] = v - p[o3];
    d[4] = v - p[o4]; d[5] = v - p[o5]; d[6] = v - p[o6]; d[7] = v - p[o7];
    d[8] = v - p[o8]; d[9] = v - p[o9]; d[10] = v - p[o10]; d[11] = v - p[o11];
    d[12] = v - p[o12]; d[13] = v - p[o13]; d[14] = v - p[o14]; d[15] = v - p[o15];
    const int m = arc_value(d);
    if (m > tq) sM[(yy + 1) * kStride + xx + 1] = (u8)m;
  }
  __syncthreads();
  // pass 3: strict 3x3 maxima with m >= 2; compact them to the front of sList (as index | m is
  // re-read from sM).  Reuse of sList is safe: every thread finishes reading entry i before any
  // thread can overwrite it only if writes go to a second region, so peaks are written behind n.
  unsigned short* sPk = sList + n;  // n + peaks <= 2 * (iw*ih) / ... bounded below
  for (int i = tid; i < n; i += 256) {
    const int e = sList[i], yy = e >> 8, xx = e & 255;
    const u8* q = sM + (yy + 1) * kStride + xx + 1;
    const int m = q[0];
    if (m >= 2 && m > q[-1] && m > q[1] && m > q[-kStride - 1] && m > q[-kStride] && m > q[-kStride + 1] &&
        m > q[kStride - 1] && m > q[kStride] && m > q[kStride + 1]) {
      sPk[atomicAdd(&sPeaks, 1)] = (unsigned short)e;
      if (m > tIni) atomicAdd(&sIni, 1);
      if (m > tMin) atomicAdd(&sMin, 1);
    }
  }
  __syncthreads();
  const int th = sIni > 0 ? tIni : tMin;
  const int nEmit = sIni > 0 ? sIni : sMin;
  if (nEmit == 0) return;
  u32* cnt = P.candCount + (size_t)slot * kMaxLevels + c.level;
  if (tid == 0) sBase = (int)atomicAdd(cnt, (u32)nEmit);
  __syncthreads();
  const int base = sBase;
  if (base + nEmit > L.candCap) {
    if (tid == 0) atomicMin(P.status + slot, ORBX_ERR_CAPACITY);
    return;
  }
  const size_t off = (size_t)slot * P.candTotal + L.candOff + base;
  const u8* rootTab = P.rootTab + L.rootTabOff;
  const int np = sPeaks;
  for (int i = tid; i < np; i += 256) {
    const int e = sPk[i], yy = e >> 8, xx = e & 255;
    const int m = sM[(yy + 1) * kStride + xx + 1];
    if (m > th) {
      const int k = atomicAdd(&sCursor, 1);
      // ROI-relative (xx + 3, yy + 3), shifted by the cell origin: coordinates relative to (16,16)
      const int x = xx + 3 + c.sx, y = yy + 3 + c.sy;
      const u32 order = ((u32)c.index << 14) | ((u32)(yy + 3) << 7) | (u32)(xx + 3);
      P.candKey[off + k] = ((u64)(u32)(m - 1) << 32) | (u64)(~order);
      P.candXY[off + k] = (u32)x | ((u32)y << 16);
      P.candNode[off + k] = (u32)(L.nIni - 1 - __ldg(rootTab + x));  // root i lives at node nIni-1-i
    }
  }
}

int launch_fast(const Params& P, int nSlots, cudaStream_t st) {
  dim3 grid(P.nCells, nSlots);
  k_fast<<<grid, 256, 0, st>>>(P);
  return 1;
}

}  // namespace orbx
