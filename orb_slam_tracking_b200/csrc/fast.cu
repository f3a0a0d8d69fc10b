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
// Pass 1 is the hot loop: the opposing-pair rejection (every arc of 9 contains one pixel of each
// opposing ring pair, so a pair with both |d| <= t proves m <= t) runs on FOUR pixels per thread
// with byte-SWAR arithmetic on 32-bit words of the staged ROI; only the few surviving pixels get
// the exact arc value.
//
// Survivors are appended to the level's candidate list with an atomic cursor; each carries its
// position in the reference's emission order (cell row-major, then y, then x) inside its key, which
// is all the quadtree needs ("first strict maximum" == max response, then min order).
#include "common.cuh"

namespace orbx {

constexpr int kSB = 96;          // staged row stride in bytes; byte b of a row <-> level column xb - 4 + b
constexpr int kSW = kSB / 4;     // ... in words
constexpr int kMaxRows = 76;     // hCell + 6 < 76

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

// bit 0 of each byte of the result = (that byte of a) > t, with K = (255 - t) * 0x00010001
__device__ __forceinline__ u32 bytes_gt(u32 a, u32 K) {
  const u32 e = (a & 0x00ff00ffu) + K, o = ((a >> 8) & 0x00ff00ffu) + K;
  return ((e >> 8) & 0x00010001u) | (o & 0x01000100u);
}
// 4 bytes starting at byte offset 4g + dx (dx in [-3, 3]) of a staged row
template <int DX>
__device__ __forceinline__ u32 word_at(const u32* row, int g) {
  if (DX == 0) return row[g];
  if (DX > 0) return __funnelshift_r(row[g], row[g + 1], 8 * DX);
  return __funnelshift_r(row[g - 1], row[g], 32 + 8 * DX);
}

__global__ void __launch_bounds__(256) k_fast(const __grid_constant__ Params P) {
  __shared__ u32 sImgW[kMaxRows * kSW];
  __shared__ u32 sMW[kMaxRows * kSW];            // m-map, same geometry, zero outside the interior
  __shared__ unsigned short sList[kMaxRows * kSB];
  __shared__ int sN, sPeaks, sIni, sMin, sBase, sCursor;
  u8* const sImg = reinterpret_cast<u8*>(sImgW);
  u8* const sM = reinterpret_cast<u8*>(sMW);

  const CellDesc c = P.cells[blockIdx.x];
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[c.level];
  const int cw = c.x1 - c.x0, ch = c.y1 - c.y0;
  if (cw < 7 || ch < 7) return;  // cv::FAST returns nothing for such ROIs
  const int tid = threadIdx.x;
  const int a = c.x0 & 3, xb = c.x0 - a;       // aligned base column; ROI column rx <-> byte rx + 4 + a
  const int nW = (a + cw + 3) >> 2;             // words per row holding ROI bytes (staged words 1 .. nW)
  const u8* __restrict__ img = level_ptr((const u8*)P.pyr, P, slot, c.level) + (size_t)c.y0 * L.pitch + xb;

  for (int i = tid; i < ch * kSW; i += 256) {
    const int r = i / kSW, k = i - r * kSW;
    u32 v = 0;
    if (k >= 1 && k <= nW) v = __ldg(reinterpret_cast<const u32*>(img + (size_t)r * L.pitch) + (k - 1));
    sImgW[i] = v;
    sMW[i] = 0;
  }
  if (tid == 0) { sN = 0; sPeaks = 0; sIni = 0; sMin = 0; sCursor = 0; }
  __syncthreads();

  const int tIni = min(max(P.iniTh, 0), 255), tMin = min(max(P.minTh, 0), 255);
  const int tq = min(tIni, tMin);
  const u32 K = (u32)(255 - tq) * 0x00010001u;
  const int bLo = a + 4 + 3, bHi = a + 4 + cw - 3;  // interior bytes [bLo, bHi)
  const int g0 = bLo >> 2, g1 = (bHi - 1) >> 2;

  // pass 1: opposing-pair rejection on 4 pixels at a time
  for (int ry = 3 + (tid >> 4); ry < ch - 3; ry += 16) {
    const u32* r0 = sImgW + ry * kSW;
    for (int g = g0 + (tid & 15); g <= g1; g += 16) {
      // bit 0 of byte j set <=> pixel 4g + j is in the interior and not yet rejected
      u32 alive = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) alive |= (u32)((4 * g + j) >= bLo && (4 * g + j) < bHi) << (8 * j);
      const u32 v = r0[g];
#define ORBX_PAIR(DY, DX)                                                                    \
  if (alive) {                                                                                \
    const u32 pa = word_at<DX>(r0 + (DY) * kSW, g), pb = word_at<-(DX)>(r0 - (DY) * kSW, g); \
    alive &= bytes_gt(__vabsdiffu4(v, pa), K) | bytes_gt(__vabsdiffu4(v, pb), K);            \
  }
      ORBX_PAIR(3, 0) ORBX_PAIR(0, 3) ORBX_PAIR(2, 2) ORBX_PAIR(-2, 2)
      ORBX_PAIR(3, 1) ORBX_PAIR(1, 3) ORBX_PAIR(-1, 3) ORBX_PAIR(-3, 1)
#undef ORBX_PAIR
      while (alive) {
        const int j = (__ffs((int)alive) - 1) >> 3;
        alive &= ~(1u << (8 * j));
        sList[atomicAdd(&sN, 1)] = (unsigned short)((ry << 8) | (4 * g + j));
      }
    }
  }
  __syncthreads();
  const int n = sN;
  // ring offsets in staged bytes, cv::FAST order
  constexpr int o0 = 3 * kSB, o1 = 3 * kSB + 1, o2 = 2 * kSB + 2, o3 = kSB + 3, o4 = 3, o5 = -kSB + 3,
                o6 = -2 * kSB + 2, o7 = -3 * kSB + 1, o8 = -3 * kSB, o9 = -3 * kSB - 1, o10 = -2 * kSB - 2,
                o11 = -kSB - 3, o12 = -3, o13 = kSB - 3, o14 = 2 * kSB - 2, o15 = 3 * kSB - 1;
  // pass 2: exact segment-test value of the surviving pixels
  for (int i = tid; i < n; i += 256) {
    const int e = sList[i], ry = e >> 8, b = e & 255;
    const u8* p = sImg + ry * kSB + b;
    const int v = p[0];
    int d[16];
    d[0] = v - p[o0]; d[1] = v - p[o1]; d[2] = v - p[o2]; d[3] = v - p[o3];
    d[4] = v - p[o4]; d[5] = v - p[o5]; d[6] = v - p[o6]; d[7] = v - p[o7];
    d[8] = v - p[o8]; d[9] = v - p[o9]; d[10] = v - p[o10]; d[11] = v - p[o11];
    d[12] = v - p[o12]; d[13] = v - p[o13]; d[14] = v - p[o14]; d[15] = v - p[o15];
    const int m = arc_value(d);
    if (m > tq) sM[ry * kSB + b] = (u8)m;
  }
  __syncthreads();
  // pass 3: strict 3x3 maxima with m >= 2, written behind the pass-1 list
  unsigned short* sPk = sList + n;
  for (int i = tid; i < n; i += 256) {
    const int e = sList[i], ry = e >> 8, b = e & 255;
    const u8* q = sM + ry * kSB + b;
    const int m = q[0];
    if (m >= 2 && m > q[-1] && m > q[1] && m > q[-kSB - 1] && m > q[-kSB] && m > q[-kSB + 1] &&
        m > q[kSB - 1] && m > q[kSB] && m > q[kSB + 1]) {
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
    const int e = sPk[i], ry = e >> 8, b = e & 255;
    const int m = sM[ry * kSB + b];
    if (m > th) {
      const int k = atomicAdd(&sCursor, 1);
      const int rx = b - 4 - a;  // ROI-relative column; the reference shifts by the cell origin (:1134-1135)
      const int x = rx + c.sx, y = ry + c.sy;  // relative to (16,16)
      const u32 order = ((u32)c.index << 14) | ((u32)ry << 7) | (u32)rx;
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
