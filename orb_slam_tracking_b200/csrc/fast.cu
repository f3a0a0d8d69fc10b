// Per-cell FAST-9/16 with 3x3 non-maximum suppression and the min-threshold retry
// (ORBextractor::ComputeKeyPointsOctTree cell loop, Features/ORBextractor.cpp:1051-1141, calling
// cv::FAST(roi, kps, th, true) at :1109 and :1119; arithmetic in SURVEY.md appendix A3).
//
// One CTA of 128 threads per (cell, slot); the cells of ALL levels are in one table, so one launch
// covers a batch.  The cell ROI (from the 16-byte aligned column left of it) is staged into shared memory by ONE TMA
// tile copy (cp.async.bulk.tensor.3d over a per-level {x, y, slot} tensor map, completion on an
// mbarrier); out-of-image box bytes are zero-filled by the hardware and never tested.
//
// The segment-test value m(p) = max over the 16 arcs of 9 of min(v - ring) / min(ring - v) does not
// depend on the threshold: p is a corner at t iff m > t and its NMS score is m - 1.  A corner
// survives NMS at t iff m >= 2 and m is a strict 3x3 maximum of the m-map restricted to the
// cell's tested interior (a neighbour that is not a corner at t has m' <= t < m anyway).  So the
// CTA computes the peaks once and then applies the reference's rule: keep the peaks with
// m > iniTh, or, if the cell has none, the peaks with m > minTh.
//
// Pass 1 is the hot loop: the opposing-pair rejection (every arc of 9 contains one pixel of each
// opposing ring pair, so a pair with both |d| <= t proves m <= t) runs on 2 x 4 pixels per thread (two vertically
// adjacent words, which share 6 of their 8 staged rows) with byte-SWAR arithmetic on 32-bit words of the staged ROI: VABSDIFF4 per ring word, then a
// carry-free "byte > t" in bit 7 of each byte.  The four pairs that live on rows {-3, 0, +3} go
// first; only words with a pixel still alive touch the other four rows, and only the few
// surviving pixels get the exact arc value (pass 2).
//
// Survivors are appended to the level's candidate list with an atomic cursor; each carries its
// position in the reference's emission order (cell row-major, then y, then x) inside its key, which
// is all the quadtree needs ("first strict maximum" == max response, then min order).
#include "common.cuh"

namespace orbx {

constexpr int kFastThreads = 128;

// m(p) for one pixel.  Both polarities ride in one register per ring position: e_k = v - p_k + 256 in the
// low half-word and 512 - e_k = p_k - v + 256 in the high one, so the min-of-9 / max-of-16 ladder below
// (VIMNMX.U16x2) yields A + 256 and B + 256 together.
__device__ __forceinline__ int arc_value(const u32 (&x)[16]) {
  // min over 9 consecutive ring positions = min3 of three min3 (DPX VIMNMX3.U16x2), then a max3 tree over the 16 arcs
  u32 mn3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) mn3[k] = __vimin3_u16x2(x[k], x[(k + 1) & 15], x[(k + 2) & 15]);
  u32 a[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) a[k] = __vimin3_u16x2(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
  u32 b[6];
#pragma unroll
  for (int k = 0; k < 5; ++k) b[k] = __vimax3_u16x2(a[3 * k], a[3 * k + 1], a[3 * k + 2]);
  b[5] = a[15];
  const u32 best = __vimax3_u16x2(__vimax3_u16x2(b[0], b[1], b[2]), __vimax3_u16x2(b[3], b[4], b[5]), 0u);
  return (int)max(best & 0xffffu, best >> 16) - 256;
}
// base = (256 - v) << 16 | (v + 256); adding p * 0xFFFF moves p from the low half to the high one: one multiply-add
__device__ __forceinline__ u32 ring_pack(u32 base, u32 p) { return p * 0xFFFFu + base; }

// bit 7 of each byte of the result = (that byte of a) > t  OR  (that byte of b) > t.
//   t < 128:  K = (127 - t) * 0x01010101; (x & 0x7f) + K carries into bit 7 iff (x & 0x7f) > t, and x >= 128 > t.
//   t >= 128: K = (255 - t) * 0x01010101; x > t iff bit 7 of x is set and (x & 0x7f) + K carries.
// MODE 0: t < 128 and t + 1 a power of two (the defaults 7 and 0): a <= t and b <= t  <=>  (a | b) <= t, one compare for both
// MODE 1: any t < 128;  MODE 2: t >= 128
template <int MODE>
__device__ __forceinline__ u32 any_gt(u32 a, u32 b, u32 K) {
  if (MODE == 0) {
    const u32 x = a | b;
    return ((x & 0x7f7f7f7fu) + K) | x;
  }
  const u32 sa = (a & 0x7f7f7f7fu) + K, sb = (b & 0x7f7f7f7fu) + K;
  if (MODE == 2) return (sa & a) | (sb & b);
  return sa | sb | a | b;
}
// 4 bytes starting at byte offset 4g + DX (DX in [-3, 3]) given the words g-1, g, g+1 of a staged row
template <int DX>
__device__ __forceinline__ u32 word_at(u32 wl, u32 wc, u32 wr) {
  if (DX == 0) return wc;
  if (DX > 0) return __funnelshift_r(wc, wr, 8 * DX);
  return __funnelshift_r(wl, wc, 32 + 8 * DX);
}

__device__ __forceinline__ u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }

struct Row3 { u32 l, c, r; };  // words g-1, g, g+1 of one staged row
__device__ __forceinline__ Row3 load3(const u32* __restrict__ row, int g) { return Row3{row[g - 1], row[g], row[g + 1]}; }

// pairs on rows {-3, 0, +3}: (dy,dx) = (3,0), (0,3), (3,1), (-3,1) and their opposites
template <int HI>
__device__ __forceinline__ u32 pairs_rows_0_3(u32 alive, const Row3& m, const Row3& c, const Row3& p, u32 K) {
  const u32 cc = c.c;
  alive &= any_gt<HI>(__vabsdiffu4(cc, p.c), __vabsdiffu4(cc, m.c), K);
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<3>(c.l, c.c, c.r)), __vabsdiffu4(cc, word_at<-3>(c.l, c.c, c.r)), K);
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<1>(p.l, p.c, p.r)), __vabsdiffu4(cc, word_at<-1>(m.l, m.c, m.r)), K);
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<1>(m.l, m.c, m.r)), __vabsdiffu4(cc, word_at<-1>(p.l, p.c, p.r)), K);
  return alive;
}
// pairs on rows +-2: (2,2)/(-2,-2), (-2,2)/(2,-2) and rows +-1: (1,3)/(-1,-3), (-1,3)/(1,-3)
template <int HI>
__device__ __forceinline__ u32 pairs_rows_1_2(u32 alive, u32 cc, const Row3& b2, const Row3& b1, const Row3& a1, const Row3& a2, u32 K) {
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<2>(a2.l, a2.c, a2.r)), __vabsdiffu4(cc, word_at<-2>(b2.l, b2.c, b2.r)), K);
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<2>(b2.l, b2.c, b2.r)), __vabsdiffu4(cc, word_at<-2>(a2.l, a2.c, a2.r)), K);
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<3>(a1.l, a1.c, a1.r)), __vabsdiffu4(cc, word_at<-3>(b1.l, b1.c, b1.r)), K);
  alive &= any_gt<HI>(__vabsdiffu4(cc, word_at<3>(b1.l, b1.c, b1.r)), __vabsdiffu4(cc, word_at<-3>(a1.l, a1.c, a1.r)), K);
  return alive;
}

// Two vertically adjacent words (rows y and y+1 of the same word column) per call: the 8 staged rows y-3 .. y+4
// serve both.  r0 points at row y.  The pairs of rows {-3,0,+3} go first; the other four rows are only
// evaluated for a word that still has a pixel alive.
template <int HI>
__device__ __forceinline__ void pass1_words(const u32* __restrict__ r0, int SW, int g, u32& aliveA, u32& aliveB, u32 K) {
  const Row3 rm3 = load3(r0 - 3 * SW, g), r00 = load3(r0, g), rp3 = load3(r0 + 3 * SW, g);
  const Row3 rm2 = load3(r0 - 2 * SW, g), rp1 = load3(r0 + SW, g), rp4 = load3(r0 + 4 * SW, g);
  aliveA = pairs_rows_0_3<HI>(aliveA, rm3, r00, rp3, K);
  aliveB = pairs_rows_0_3<HI>(aliveB, rm2, rp1, rp4, K);
  if (aliveA | aliveB) {
    const Row3 rm1 = load3(r0 - SW, g), rp2 = load3(r0 + 2 * SW, g);
    if (aliveA) aliveA = pairs_rows_1_2<HI>(aliveA, r00.c, rm2, rm1, rp1, rp2, K);
    if (aliveB) aliveB = pairs_rows_1_2<HI>(aliveB, rp1.c, rm1, r00, rp2, rp3, K);
  }
}

__global__ void __launch_bounds__(kFastThreads) k_fast(const __grid_constant__ Params P, const __grid_constant__ TmaMaps M) {
  extern __shared__ __align__(128) u8 smem[];
  __shared__ __align__(8) u64 sBar;
  __shared__ int sN, sPeaks, sIni, sMin, sBase, sCursor;

  const CellDesc c = P.cells[blockIdx.x];
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[c.level];
  const int cw = c.x1 - c.x0, ch = c.y1 - c.y0;
  const int tid = threadIdx.x;
  const int SB = P.fastBoxW, SW = SB >> 2;  // staged row stride in bytes / words
  // TMA tile origins must be 16-byte aligned in the innermost dimension: the tile starts at the aligned column
  // xs <= x0 - 1, so byte b of a staged row <-> level column xs + b and ROI column rx <-> byte rx + off, off in [1, 16]
  const int xs = (c.x0 - 1) & ~15, off = c.x0 - xs;
  const int tileBytes = SB * P.fastBoxH;
  u8* const sImg = smem;
  u8* const sM = smem + tileBytes;            // m-map, same geometry, zero outside the interior
  unsigned short* const sList = reinterpret_cast<unsigned short*>(smem + 2 * tileBytes);
  const u32* const sImgW = reinterpret_cast<const u32*>(sImg);

  if (tid == 0) {
    sN = 0; sPeaks = 0; sIni = 0; sMin = 0; sCursor = 0;
    const u32 bar = smem_addr(&sBar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(tileBytes) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_addr(sImg)), "l"(reinterpret_cast<u64>(&M.lvl[c.level])), "r"(xs), "r"((int)c.y0), "r"(slot + P.slot0), "r"(bar)
        : "memory");
  }
  {  // zero the m-map while the tile is in flight
    uint4* z = reinterpret_cast<uint4*>(sM);
    for (int i = tid; i < (tileBytes >> 4); i += kFastThreads) z[i] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();  // sBar initialised (and the zeroes visible) before anyone polls
  {
    const u32 bar = smem_addr(&sBar);
    u32 done = 0;
    while (!done) {
      asm volatile(
          "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}\n"
          : "=r"(done) : "r"(bar) : "memory");
    }
  }

  const int tIni = min(max(P.iniTh, 0), 255), tMin = min(max(P.minTh, 0), 255);
  const int tq = min(tIni, tMin);
  const int bLo = off + 3, bHi = off + cw - 3;  // interior bytes [bLo, bHi)
  const int g0 = bLo >> 2, g1 = (bHi - 1) >> 2;
  const int nW = g1 - g0 + 1, nR = ch - 6;

  // pass 1: opposing-pair rejection on 4 pixels at a time, items = (interior row, word) flattened over the CTA
  {
    const u32 K = (u32)(tq < 128 ? 127 - tq : 255 - tq) * 0x01010101u;
    // items = (pair of interior rows, word): thread-private indices advance by the CTA size, row-major
    const int nRP = (nR + 1) >> 1;
    const int stepR = kFastThreads / nW, stepG = kFastThreads - stepR * nW;
    const int nIter = (nRP * nW + kFastThreads - 1) / kFastThreads;  // the same trip count for every lane: warp collectives below
    const int lane = tid & 31;
    int rp = tid / nW, g = tid - rp * nW;
    for (int it = 0; it < nIter; ++it, rp += stepR, g += stepG) {
      if (g >= nW) { g -= nW; ++rp; }
      u32 aliveA = 0, aliveB = 0;
      const int gw = g0 + g, ry = 2 * rp;
      if (ry < nR) {
        u32 m = 0x80808080u;
        if (4 * gw < bLo) m &= 0xffffffffu << (8 * (bLo - 4 * gw));
        if (4 * gw + 4 > bHi) m &= 0xffffffffu >> (8 * (4 * gw + 4 - bHi));
        aliveA = m;
        aliveB = ry + 1 < nR ? m : 0u;  // the row below the last interior row is staged (apron) but not tested
        const u32* r0 = sImgW + (ry + 3) * SW;
        if (tq >= 128) pass1_words<2>(r0, SW, gw, aliveA, aliveB, K);
        else if ((tq & (tq + 1)) == 0) pass1_words<0>(r0, SW, gw, aliveA, aliveB, K);
        else pass1_words<1>(r0, SW, gw, aliveA, aliveB, K);
      }
      // append: the list order is irrelevant (pass 2 / 3 take any order, the emission carries its own order key), so a
      // lane with survivors just reserves its slots
      const int cnt = __popc(aliveA) + __popc(aliveB);
      if (cnt == 0) continue;
      int base = atomicAdd(&sN, cnt);
      while (aliveA) {
        const int j = (__ffs((int)aliveA) - 1) >> 3;
        aliveA &= aliveA - 1;
        sList[base++] = (unsigned short)(((ry + 3) << 8) | (4 * gw + j));
      }
      while (aliveB) {
        const int j = (__ffs((int)aliveB) - 1) >> 3;
        aliveB &= aliveB - 1;
        sList[base++] = (unsigned short)(((ry + 4) << 8) | (4 * gw + j));
      }
    }
  }
  __syncthreads();
  const int n = sN;
  // pass 2: exact segment-test value of the surviving pixels (ring offsets in cv::FAST order)
  for (int i = tid; i < n; i += kFastThreads) {
    const int e = sList[i], ry = e >> 8, b = e & 255;
    const u8* p = sImg + ry * SB + b;
    const u32 vb = ((256u - (u32)p[0]) << 16) | ((u32)p[0] + 256u);
    u32 x[16];
    x[0] = ring_pack(vb, p[3 * SB]); x[1] = ring_pack(vb, p[3 * SB + 1]); x[2] = ring_pack(vb, p[2 * SB + 2]); x[3] = ring_pack(vb, p[SB + 3]);
    x[4] = ring_pack(vb, p[3]); x[5] = ring_pack(vb, p[-SB + 3]); x[6] = ring_pack(vb, p[-2 * SB + 2]); x[7] = ring_pack(vb, p[-3 * SB + 1]);
    x[8] = ring_pack(vb, p[-3 * SB]); x[9] = ring_pack(vb, p[-3 * SB - 1]); x[10] = ring_pack(vb, p[-2 * SB - 2]); x[11] = ring_pack(vb, p[-SB - 3]);
    x[12] = ring_pack(vb, p[-3]); x[13] = ring_pack(vb, p[SB - 3]); x[14] = ring_pack(vb, p[2 * SB - 2]); x[15] = ring_pack(vb, p[3 * SB - 1]);
    const int m = arc_value(x);
    if (m > tq) sM[ry * SB + b] = (u8)m;
  }
  __syncthreads();
  // pass 3: strict 3x3 maxima with m >= 2, written behind the pass-1 list
  unsigned short* sPk = sList + n;
  for (int i = tid; i < n; i += kFastThreads) {
    const int e = sList[i], ry = e >> 8, b = e & 255;
    const u8* q = sM + ry * SB + b;
    const int m = q[0];
    if (m >= 2 && m > q[-1] && m > q[1] && m > q[-SB - 1] && m > q[-SB] && m > q[-SB + 1] &&
        m > q[SB - 1] && m > q[SB] && m > q[SB + 1]) {
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
  const size_t co = (size_t)slot * P.candTotal + L.candOff + base;
  const u8* rootTab = P.rootTab + L.rootTabOff;
  const int np = sPeaks;
  for (int i = tid; i < np; i += kFastThreads) {
    const int e = sPk[i], ry = e >> 8, b = e & 255;
    const int m = sM[ry * SB + b];
    if (m > th) {
      const int k = atomicAdd(&sCursor, 1);
      const int rx = b - off;  // ROI-relative column; the reference shifts by the cell origin (:1134-1135)
      const int x = rx + c.sx, y = ry + c.sy;  // relative to (16,16)
      const u32 order = ((u32)c.index << 14) | ((u32)ry << 7) | (u32)rx;
      P.candKey[co + k] = ((u64)(u32)(m - 1) << 32) | (u64)(~order);
      P.candXY[co + k] = (u32)x | ((u32)y << 16);
      P.candNode[co + k] = (u32)(L.nIni - 1 - __ldg(rootTab + x));  // root i lives at node nIni-1-i
    }
  }
}

int launch_fast(const Params& P, const TmaMaps& M, int nSlots, cudaStream_t st) {
  dim3 grid(P.nCells, nSlots);
  // image tile + m-map + pass-1 list and peak list (<= 1.25 entries per staged byte, u16 each)
  const size_t smem = (size_t)P.fastBoxW * P.fastBoxH * 9 / 2 + 64;
  k_fast<<<grid, kFastThreads, smem, st>>>(P, M);
  return 1;
}

}  // namespace orbx
