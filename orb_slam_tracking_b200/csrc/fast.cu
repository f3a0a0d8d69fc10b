// Per-cell FAST-9/16 with 3x3 non-maximum suppression and the min-threshold retry
// (ORBextractor::ComputeKeyPointsOctTree cell loop, Features/ORBextractor.cpp:1051-1141, calling
// cv::FAST(roi, kps, th, true) at :1109 and :1119; arithmetic in SURVEY.md appendix A3).
//
// Work item = STRIP: a run of consecutive cells of one cell row of one (level, slot); the strips of all levels are in one
// table, so one launch covers a batch.  The kernel is persistent: gridDim.x = SMs x resident CTAs, each CTA walks the
// items blockIdx.x, blockIdx.x + gridDim.x, ... with a two-stage tile ring: while it works on the tile in one buffer, the
// next item's tile is in flight into the other one (ONE cp.async.bulk.tensor.3d per tile over a per-level {x, y, slot}
// tensor map, completion on the buffer's mbarrier; rows beyond the level are zero-filled by the hardware and never tested).
//
// The segment-test value m(p) = max over the 16 arcs of 9 of min(v - ring) / min(ring - v) does not depend on the
// threshold: p is a corner at t iff m > t and its NMS score is m - 1.  A corner survives NMS at t iff m >= 2 and m is a
// strict 3x3 maximum among the corners at t of ITS cell's tested interior (everything else counts as 0).  The reference's
// schedule is kept: the whole strip runs at iniThFAST; the cells without a keypoint then run again at minThFAST (:1109-1119).
//
// One run = four data-parallel phases over compacted lists in shared memory (each phase keeps all lanes busy):
//   A   every aligned word pair (8 pixels) of the interior: the vertical opposing pair (0,+-3) only -- two byte-SWAR
//       absolute differences per word (VABSDIFF4) and a carry-free "byte > t" -- needs nothing but aligned 64-bit loads;
//       words with a pixel alive go to the word list (warp-aggregated append).  Flat image areas end here.
//   BC  one thread per listed word: all 8 opposing pairs (every arc of 9 contains one pixel of each pair, so a pair with
//       both |d| <= t proves m <= t); rows {-3, 0, +3} first, the other four rows only for words still alive; survivors
//       go to the pixel list.
//   P2  one thread per listed pixel: the exact m (both polarities in 16-bit halves, VIMNMX3 min/max ladder); m > t is
//       written to the m-map (zero everywhere else).
//   P3  one thread per listed pixel with m > t: strict 3x3 maximum inside its cell -> the keypoint is appended to the
//       level's candidate list in HBM right away, carrying its position in the reference's emission order (cell
//       row-major, then y, then x) inside its key, which is all the quadtree needs ("first strict maximum" == max
//       response, then min order); per-cell keypoint counters decide the minThFAST retry.
// The m-map entries written by a run are cleared through the pixel list afterwards (no per-strip zero fill).
#include <algorithm>

#include "common.cuh"

namespace orbx {

constexpr int kFastThreads = 32 * kFastWarps;
#ifndef ORBX_FAST_BC
#define ORBX_FAST_BC 2  // pair tests of phase BC: 2 = all six, 1 = the two on rows +-3 only, 0 = none (every pixel that passes phase A gets the exact value)
#endif
#ifndef ORBX_FAST_STAGES
#define ORBX_FAST_STAGES 2
#endif
constexpr int kFastStages = ORBX_FAST_STAGES;  // tile buffers of the ring
#ifndef ORBX_FAST_CTAS
#define ORBX_FAST_CTAS (ORBX_FAST_STAGES == 1 ? 4 : 3)
#endif
constexpr int kFastCtasPerSm = ORBX_FAST_CTAS;   // register cap: 65536 / (3 * 256) = 85

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

// Bit 7 of each byte of the result is set if that byte of a OR that byte of b is > t -- and possibly when neither is:
// the compare runs on a | b, which is >= max(a, b) and <= t whenever both are <= t and t + 1 is a power of two (the
// defaults 7 and 0).  For other thresholds a few more pixels pass the pair tests; the tests are only a filter in front of
// the exact segment-test value, whose comparison with t decides.
//   t < 128  (HI = 0): K = (127 - t) * 0x01010101; (x & 0x7f) + K carries into bit 7 iff (x & 0x7f) > t, and x >= 128 > t.
//   t >= 128 (HI = 1): K = (255 - t) * 0x01010101; x > t iff bit 7 of x is set and (x & 0x7f) + K carries.
// Three instructions, spelled out as lop3 so that they stay three: (a | b) & 0x7f.., + K, and the combination with a | b.
template <int HI>
__device__ __forceinline__ u32 any_gt(u32 a, u32 b, u32 K) {
  u32 x, r;
  asm("lop3.b32 %0, %1, %2, 0x7f7f7f7f, 0xa8;" : "=r"(x) : "r"(a), "r"(b));            // (a | b) & 0x7f7f7f7f
  const u32 s = x + K;
  if (HI) asm("lop3.b32 %0, %1, %2, %3, 0xa8;" : "=r"(r) : "r"(a), "r"(b), "r"(s));    // (a | b) & s
  else asm("lop3.b32 %0, %1, %2, %3, 0xfe;" : "=r"(r) : "r"(a), "r"(b), "r"(s));       // a | b | s
  return r;
}
// 4 bytes starting at byte offset 4g + DX (DX in [-3, 3]) given the words g-1, g, g+1 of a staged row
template <int DX>
__device__ __forceinline__ u32 word_at(u32 wl, u32 wc, u32 wr) {
  if (DX == 0) return wc;
  if (DX > 0) return __funnelshift_r(wc, wr, 8 * DX);
  return __funnelshift_r(wl, wc, 32 + 8 * DX);
}


// ballot over "x & m != 0": one LOP3 with a predicate result in front of the vote
__device__ __forceinline__ u32 ballot_bits(u32 x, u32 m) {
  u32 b;
  asm volatile("{\n .reg .pred p;\n .reg .b32 t;\n and.b32 t, %1, %2;\n setp.ne.u32 p, t, 0;\n vote.sync.ballot.b32 %0, p, 0xffffffff;\n}\n"
               : "=r"(b) : "r"(x), "r"(m));
  return b;
}

// A shared-memory pointer the compiler must keep in a register: without this it recomputes list bases such as
// warp * segment length + offset (a constant-bank load, a multiply and an add) in front of every access in the chunk loops.
template <typename T>
__device__ __forceinline__ T* keep_shared(T* p) {
  u32 a = (u32)__cvta_generic_to_shared(p);
  asm volatile("" : "+r"(a));
  return reinterpret_cast<T*>(__cvta_shared_to_generic((size_t)a));
}
__device__ __forceinline__ int keep(int v) { asm volatile("" : "+r"(v)); return v; }

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


// dynamic shared memory of k_fast: two tile buffers, the m-map, and per warp a word / peak segment (32-bit entries) and a
// pixel segment (16-bit entries)
static size_t fast_smem_bytes(const Params& P) {
  return (size_t)(kFastStages + 1) * P.fastTileMax + (size_t)kFastWarps * (4 * (size_t)P.fastSegW + 2 * (size_t)P.fastSegP);
}

struct StripCtx {
  const u32* imgW;        // staged tile as words
  u8* mm;                 // m-map, same geometry as the tile
  u32* segW;              // this warp's word segment: alive bits (7, 15, 23, 31) | row << 8 | word; reused for its peaks as 16-bit pixel entries
                          // (at most every other pixel of a row is a peak: two per word)
  unsigned short* segP;   // this warp's pixel segment: row << 9 | byte column
  int* cellIni;           // per cell of the strip: set by a keypoint of the iniThFAST run
  int SB, SW;             // staged row stride in bytes / words
  int bLo, bHi;           // interior byte columns [bLo, bHi) of a staged row
  int nR;                 // interior rows: tile rows [3, 3 + nR)
  int wCell;
  u32 cellMul;
  // emission
  int slot, level, sx, sy, index0, nIniM1, candCap;
  u32 candOff, rootTabOff;
  int retryMax;           // a cell with at most this many keypoints at iniThFAST is tested again at minThFAST (and its iniThFAST keypoints dropped)
  int orderBits;          // bits per coordinate of the emission-order key
};

// bytes of word g (bit 7 each) that lie inside [bLo, bHi): the low max(bLo - 4g, 0) and the high max(4g + 4 - bHi, 0) bytes are cut
// off by two shifts (PTX shifts by 32 or more give 0, which is what a word outside the interval needs)
__device__ __forceinline__ u32 interior_mask(int g, int bLo, int bHi) {
  u32 lo, hi;
  asm("shl.b32 %0, %1, %2;" : "=r"(lo) : "r"(0x80808080u), "r"((u32)max(8 * bLo - 32 * g, 0)));
  asm("shr.u32 %0, %1, %2;" : "=r"(hi) : "r"(0x80808080u), "r"((u32)max(32 * g + 32 - 8 * bHi, 0)));
  return lo & hi;
}
// bytes of word g (any bits) whose cell is in `cells` (retry run); a word spans at most two cells
__device__ __forceinline__ u32 cell_mask(const StripCtx& S, int g, u32 cells) {
  const int xa = max(4 * g - S.bLo, 0), xb = max(4 * g + 3 - S.bLo, 0);
  const int ja = (int)(((u32)xa * S.cellMul) >> kCellShift), jb = (int)(((u32)xb * S.cellMul) >> kCellShift);
  const u32 ka = (cells >> ja) & 1 ? 0xffffffffu : 0u, kb = (cells >> jb) & 1 ? 0xffffffffu : 0u;
  if (ja == jb) return ka;
  const int nb = jb * S.wCell + S.bLo - 4 * g;  // first byte of the word that belongs to cell jb, 1..3
  const u32 lo = 0xffffffffu >> (8 * (4 - nb));
  return (ka & lo) | (kb & ~lo);
}

// One pass of the phases over the strip at threshold t.  `cells`: the cells to test (all of them in the iniThFAST run).
// The lists are per warp: a warp filters its own rows (A), runs the remaining pair tests on its own word segment (BC) and
// the exact value on its own pixel segment (P2) without a CTA barrier; the rows of a warp are interleaved over the strip
// (r = warp, warp + 8, ...), which evens out the segment lengths.  CTA barriers only around the m-map's neighbour reads.
template <int MODE>
__device__ __forceinline__ void fast_run(const Params& P, const StripCtx& S, const int t, const bool RETRY, const u32 cells) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u32 lt = (1u << lane) - 1;
  const int SW = S.SW, SB = S.SB;
  const u32 K = (u32)(t < 128 ? 127 - t : 255 - t) * 0x01010101u;

  // ---- phase A: a warp per interior row, a lane per aligned word pair.  The strips of the default grid are at most 32 pairs
  // wide (one pass); the wide cells of ComputeKeyPointsOld take one pass per block of 32 pairs. ----
  int nW = 0;
  {
    const int gBase = (S.bLo >> 2) & ~1;
    const int nQ = ((((S.bHi - 1) >> 2) - gBase) >> 1) + 1;
    for (int qb = 0; qb < nQ; qb += 32) {
      const int g = gBase + 2 * (qb + lane);  // first word of this lane's pair
      u32 m0 = interior_mask(g, S.bLo, S.bHi), m1 = interior_mask(g + 1, S.bLo, S.bHi);
      if (RETRY) { m0 &= cell_mask(S, g, cells); m1 &= cell_mask(S, g + 1, cells); }
      const bool act = (m0 | m1) != 0;
      // one pointer and one list-entry stem walk down the warp's rows: e = row << 8 | word (a word index is below 128)
      const u32* p = S.imgW + (warp + 3) * SW + g;
      const int pStep = keep(kFastWarps * SW), o3 = keep(3 * SW);
      u32 e = ((u32)(warp + 3) << 8) + (u32)g;
      // the two pair tests of one word pair of one row (loads and arithmetic only: two rows are in flight per trip of the loop)
      auto tests = [&](const u32* pp, u32& a0, u32& a1) {
        const uint2 c = *reinterpret_cast<const uint2*>(pp);
        const uint2 n = *reinterpret_cast<const uint2*>(pp - o3);
        const uint2 s = *reinterpret_cast<const uint2*>(pp + o3);
        const u32 l = pp[-1], q = pp[2];
        a0 = m0 & any_gt<MODE>(__vabsdiffu4(c.x, n.x), __vabsdiffu4(c.x, s.x), K);
        a1 = m1 & any_gt<MODE>(__vabsdiffu4(c.y, n.y), __vabsdiffu4(c.y, s.y), K);
        a0 &= any_gt<MODE>(__vabsdiffu4(c.x, word_at<3>(l, c.x, c.y)), __vabsdiffu4(c.x, word_at<-3>(l, c.x, c.y)), K);
        a1 &= any_gt<MODE>(__vabsdiffu4(c.y, word_at<3>(c.x, c.y, q)), __vabsdiffu4(c.y, word_at<-3>(c.x, c.y, q)), K);
      };
      auto append = [&](u32 a0, u32 a1, u32 ee) {
        const u32 b0 = __ballot_sync(0xffffffffu, a0 != 0), b1 = __ballot_sync(0xffffffffu, a1 != 0);
        if (a0) S.segW[nW + __popc(b0 & lt)] = a0 | ee;
        nW += __popc(b0);
        if (a1) S.segW[nW + __popc(b1 & lt)] = a1 | (ee + 1);
        nW += __popc(b1);
      };
      int r = warp;
      for (; r + kFastWarps < S.nR; r += 2 * kFastWarps, p += 2 * pStep, e += (u32)kFastWarps << 9) {
        u32 a0 = 0, a1 = 0, c0 = 0, c1 = 0;
        if (act) { tests(p, a0, a1); tests(p + pStep, c0, c1); }
        append(a0, a1, e);
        append(c0, c1, e + ((u32)kFastWarps << 8));
      }
      if (r < S.nR) {
        u32 a0 = 0, a1 = 0;
        if (act) tests(p, a0, a1);
        append(a0, a1, e);
      }
    }
  }
  __syncwarp();
  // ---- phase BC: a lane per listed word of this warp, the other six opposing pairs ----
  int nP = 0;
  for (int i0 = 0; i0 < nW; i0 += 32) {
    const int i = i0 + lane;
    u32 alive = 0, e = 0;
    if (i < nW) {
      e = S.segW[i];
      alive = e & 0x80808080u;
      const int g = e & 0x7f;
      const u32* r0 = S.imgW + ((e >> 8) & 0x7f) * SW;
#if ORBX_FAST_BC >= 1
      const Row3 rm3 = load3(r0 - 3 * SW, g), r00 = load3(r0, g), rp3 = load3(r0 + 3 * SW, g);
      const u32 cc = r00.c;
      alive &= any_gt<MODE>(__vabsdiffu4(cc, word_at<1>(rp3.l, rp3.c, rp3.r)), __vabsdiffu4(cc, word_at<-1>(rm3.l, rm3.c, rm3.r)), K);
      alive &= any_gt<MODE>(__vabsdiffu4(cc, word_at<1>(rm3.l, rm3.c, rm3.r)), __vabsdiffu4(cc, word_at<-1>(rp3.l, rp3.c, rp3.r)), K);
#endif
      if (ORBX_FAST_BC >= 2 && alive) {
#if ORBX_FAST_BC >= 2
        const Row3 rm2 = load3(r0 - 2 * SW, g), rm1 = load3(r0 - SW, g), rp1 = load3(r0 + SW, g), rp2 = load3(r0 + 2 * SW, g);
        alive = pairs_rows_1_2<MODE>(alive, cc, rm2, rm1, rp1, rp2, K);
#endif
      }
    }
    __syncwarp();
    // append the alive pixels: a lane's pixels (0..4) take consecutive entries; the lane offsets are the prefix sum of the
    // counts, from three ballots over the bits of the count
    const int c = __popc(alive);  // alive holds bits 7, 15, 23, 31 only
    const u32 c0 = ballot_bits((u32)c, 1u), c1 = ballot_bits((u32)c, 2u), c2 = ballot_bits((u32)c, 4u);
    unsigned short* q = S.segP + (nP + __popc(c0 & lt) + 2 * __popc(c1 & lt) + 4 * __popc(c2 & lt));
    nP += __popc(c0) + 2 * __popc(c1) + 4 * __popc(c2);
    const u32 pe = ((e & 0x7f00u) << 1) | ((e & 0x7fu) << 2);  // row << 9 | 4 * word
    if (alive & 0x80u) *q++ = (unsigned short)pe;
    if (alive & 0x8000u) *q++ = (unsigned short)(pe + 1);
    if (alive & 0x800000u) *q++ = (unsigned short)(pe + 2);
    if (alive & 0x80000000u) *q = (unsigned short)(pe + 3);
  }
  __syncwarp();
  // ---- phase P2: exact segment-test value of the surviving pixels (ring offsets in cv::FAST order).  The pixel list is
  // compacted IN PLACE to the corners (m > t; a chunk's survivors land at or before the chunk): only they own an m-map entry,
  // so only they take part in the maximum search and in the clearing ----
  const u8* sImg = reinterpret_cast<const u8*>(S.imgW);
  const int SB1 = keep(SB), SB2 = keep(2 * SB), SB3 = keep(3 * SB);  // row offsets of the ring, held in registers over the chunk loop
  int nC = 0;
  for (int i0 = 0; i0 < nP; i0 += 32) {
    const int i = i0 + lane;
    bool corner = false;
    unsigned short e16 = 0;
    if (i < nP) {
      e16 = S.segP[i];
      const int e = e16, ry = e >> 9, b = e & 511;
      const u8* p = sImg + ry * SB + b;
      const u32 vb = ((256u - (u32)p[0]) << 16) | ((u32)p[0] + 256u);
      u32 x[16];
      const u8 *pa1 = p + SB1, *pa2 = p + SB2, *pa3 = p + SB3, *pb1 = p - SB1, *pb2 = p - SB2, *pb3 = p - SB3;
      x[0] = ring_pack(vb, pa3[0]); x[1] = ring_pack(vb, pa3[1]); x[2] = ring_pack(vb, pa2[2]); x[3] = ring_pack(vb, pa1[3]);
      x[4] = ring_pack(vb, p[3]); x[5] = ring_pack(vb, pb1[3]); x[6] = ring_pack(vb, pb2[2]); x[7] = ring_pack(vb, pb3[1]);
      x[8] = ring_pack(vb, pb3[0]); x[9] = ring_pack(vb, pb3[-1]); x[10] = ring_pack(vb, pb2[-2]); x[11] = ring_pack(vb, pb1[-3]);
      x[12] = ring_pack(vb, p[-3]); x[13] = ring_pack(vb, pa1[-3]); x[14] = ring_pack(vb, pa2[-2]); x[15] = ring_pack(vb, pa3[-1]);
      const int m = arc_value(x);
      corner = m > t;
      if (corner) S.mm[ry * SB + b] = (u8)m;
    }
    const u32 bc = __ballot_sync(0xffffffffu, corner);
    __syncwarp();
    if (corner) S.segP[nC + __popc(bc & lt)] = e16;
    nC += __popc(bc);
    __syncwarp();
  }
  nP = nC;
  __syncthreads();
  // ---- phase P3: strict 3x3 maxima inside the pixel's own cell (branch-free: 60 % of the listed pixels are corners) ----
  int nK = 0;
  unsigned short* const segK = reinterpret_cast<unsigned short*>(S.segW);
  for (int i0 = 0; i0 < nP; i0 += 32) {
    const int i = i0 + lane;
    bool peak = false;
    int e = 0;
    if (i < nP) {
      e = S.segP[i];
      const int ry = e >> 9, b = e & 511;
      const u8* q = S.mm + ry * SB + b;
      const int m = q[0];
      const int xb = b - S.bLo;
      const int j = (int)(((u32)xb * S.cellMul) >> kCellShift), cs = j * S.wCell;
      // the columns left of the cell's first and right of its last interior column belong to another cell: they count as 0
      const int mL = xb != cs ? max(max((int)q[-SB - 1], (int)q[-1]), (int)q[SB - 1]) : 0;
      const int mR = xb + 1 != cs + S.wCell ? max(max((int)q[-SB + 1], (int)q[1]), (int)q[SB + 1]) : 0;
      const int mN = max(max((int)q[-SB], (int)q[SB]), max(mL, mR));
      peak = m >= 2 && m > mN;
      if (peak && !RETRY) {
        if (S.retryMax == 0) S.cellIni[j] = 1;     // ComputeKeyPointsOctTree only asks "any keypoint?" (:1116)
        else atomicAdd(&S.cellIni[j], 1);          // ComputeKeyPointsOld counts them (:1343)
      }
    }
    const u32 bp = __ballot_sync(0xffffffffu, peak);
    if (peak) segK[nK + __popc(bp & lt)] = (unsigned short)e;
    nK += __popc(bp);
  }
  __syncwarp();
  // ---- phase P4: the keypoints of this run go to the level's candidate list ----
  // ComputeKeyPointsOld drops the iniThFAST keypoints of a cell that has at most retryMax of them: its emission waits for
  // the cell counts of all warps (after the barrier) and compacts the warp's peak list first.
  // The warp's keypoints take consecutive slots of the candidate list: ONE global atomic per warp.  It is issued before the
  // barrier and its result is used after it, so the round trip to L2 (about a microsecond) runs beside the barrier wait; the
  // keypoints are written after the barrier -- a warp's peaks come from its own pixel list, whose m-map entries only this warp
  // clears (below), so they are still in place.
  const bool deferred = S.retryMax > 0 && !RETRY;
  u32 k0 = 0;
  auto reserve = [&]() {
    if (nK > 0 && lane == 0) k0 = atomicAdd(P.candCount + (size_t)S.slot * kMaxLevels + S.level, (u32)nK);
  };
  if (!deferred) reserve();
  __syncthreads();  // every neighbour read of the m-map is done
  if (deferred) {
    int kept = 0;
    for (int i0 = 0; i0 < nK; i0 += 32) {
      const int i = i0 + lane;
      bool on = false;
      unsigned short e = 0;
      if (i < nK) {
        e = segK[i];
        const int xb = (int)(e & 511) - S.bLo;
        on = S.cellIni[(int)(((u32)xb * S.cellMul) >> kCellShift)] > S.retryMax;
      }
      const u32 bo = __ballot_sync(0xffffffffu, on);
      __syncwarp();
      if (on) segK[kept + __popc(bo & lt)] = e;   // kept + rank <= i: in-place compaction, a chunk at a time
      kept += __popc(bo);
      __syncwarp();
    }
    nK = kept;
    reserve();
  }
  if (nK > 0) {
    const size_t co = (size_t)S.slot * P.candTotal + S.candOff;
    const u8* rootTab = P.rootTab + S.rootTabOff;
    k0 = __shfl_sync(0xffffffffu, k0, 0);
    if (k0 + (u32)nK > (u32)S.candCap) {
      if (lane == 0) atomicMin(P.status + S.slot, ORBX_ERR_CAPACITY);
    } else {
      for (int i = lane; i < nK; i += 32) {
        const u32 pk = segK[i], k = k0 + (u32)i;
        const int ry = (int)(pk >> 9), b = (int)(pk & 511), xb = b - S.bLo;
        const int m = S.mm[ry * SB + b];
        const int j = (int)(((u32)xb * S.cellMul) >> kCellShift);
        const int rx = xb - j * S.wCell + 3;  // ROI-relative column in the pixel's cell; the reference shifts by the cell origin (:1134-1135)
        const int x = xb + 3 + S.sx, y = ry + S.sy;  // relative to (16,16)
        const u32 order = ((u32)(S.index0 + j) << (2 * S.orderBits)) | ((u32)ry << S.orderBits) | (u32)rx;
        P.candKey[co + k] = ((u64)(u32)(m - 1) << 32) | (u64)(~order);
        P.candXY[co + k] = (u32)x | ((u32)y << 16);
        P.candNode[co + k] = (u32)(S.nIniM1 - __ldg(rootTab + x));  // root i lives at node nIni-1-i
      }
    }
  }
  __syncwarp();
  for (int i = lane; i < nP; i += 32) {  // the m-map entries of this run are cleared
    const int e = S.segP[i];
    S.mm[(e >> 9) * SB + (e & 511)] = 0;
  }
}

__device__ __forceinline__ void fast_run_mode(const Params& P, const StripCtx& S, int t, bool retry, u32 cells) {
  if (t >= 128) fast_run<1>(P, S, t, retry, cells);
  else fast_run<0>(P, S, t, retry, cells);
}

__global__ void __launch_bounds__(kFastThreads, kFastCtasPerSm) k_fast(const __grid_constant__ Params P, const __grid_constant__ TmaMaps M, const int nItems) {
  extern __shared__ __align__(128) u8 smem[];
  __shared__ __align__(8) u64 sBar[2];
  __shared__ int sCellIni[32];
  const int tid = threadIdx.x;
  u8* const sMm = keep_shared(smem + kFastStages * P.fastTileMax);
  u32* const sSegW = keep_shared(reinterpret_cast<u32*>(sMm + P.fastTileMax) + (tid >> 5) * P.fastSegW);
  unsigned short* const sSegP =
      keep_shared(reinterpret_cast<unsigned short*>(reinterpret_cast<u32*>(sMm + P.fastTileMax) + kFastWarps * P.fastSegW) + (tid >> 5) * P.fastSegP);

  {
    uint4* z = reinterpret_cast<uint4*>(sMm);  // the m-map starts out zero and every run leaves it zero
    for (int i = tid; i < (P.fastTileMax >> 4); i += kFastThreads) z[i] = make_uint4(0, 0, 0, 0);
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sBar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sBar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  // The items of this CTA are blockIdx.x, blockIdx.x + gridDim.x, ...; item = slot * nStrips + strip, walked without divisions.
  const int stepSlot = (int)gridDim.x / P.nStrips, stepStrip = (int)gridDim.x - stepSlot * P.nStrips;
  int item = blockIdx.x, slot = item / P.nStrips, strip = item - slot * P.nStrips;
  auto load_desc = [&](int st) {
    union { uint4 q[3]; StripDesc d; } u;
    const uint4* g = reinterpret_cast<const uint4*>(P.strips + st);
    u.q[0] = __ldg(g); u.q[1] = __ldg(g + 1); u.q[2] = __ldg(g + 2);
    return u.d;
  };
  // thread 0: one TMA tile copy of a strip of a slot into ring buffer `buf`
  auto issue = [&](int sl, const StripDesc& d, int buf) {
    const u32 bar = smem_addr(&sBar[buf]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(d.tileBytes & 0xffffff) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_addr(smem + buf * P.fastTileMax)), "l"(reinterpret_cast<u64>(&M.lvl[d.level])), "r"((int)d.tmaX), "r"((int)d.y0),
          "r"(sl + P.slot0), "r"(bar)
        : "memory");
  };

  const int tIni = min(max(P.iniTh, 0), 255), tMin = min(max(P.minTh, 0), 255);
  if (tid == 0 && item < nItems) issue(slot, load_desc(strip), 0);
  for (int k = 0; item < nItems; ++k) {
    const int buf = kFastStages == 1 ? 0 : k & 1;
    int nslot = slot + stepSlot, nstrip = strip + stepStrip;
    if (nstrip >= P.nStrips) { nstrip -= P.nStrips; ++nslot; }
    // the other buffer was last read during item k - 1, which every thread has left (barrier at the end of the loop body)
    if (kFastStages > 1 && tid == 0 && item + (int)gridDim.x < nItems) issue(nslot, load_desc(nstrip), buf ^ 1);
    if (tid < 32) sCellIni[tid] = 0;
    const StripDesc d = load_desc(strip);
    StripCtx S;
    S.imgW = keep_shared(reinterpret_cast<const u32*>(smem + buf * P.fastTileMax));
    S.mm = sMm; S.segW = sSegW; S.segP = sSegP; S.cellIni = sCellIni;
    S.SB = d.SB; S.SW = S.SB >> 2;
    S.bLo = d.bLo; S.bHi = d.bHi; S.nR = d.nR;
    S.wCell = d.wCell; S.cellMul = d.cellMul;
    S.slot = slot; S.level = d.level; S.sx = d.sx; S.sy = d.sy; S.index0 = d.index0;
    S.nIniM1 = d.nIniM1; S.candCap = d.candCap; S.candOff = d.candOff; S.rootTabOff = d.rootTabOff;
    S.retryMax = (d.tileBytes >> 24) & 15; S.orderBits = (int)((u32)d.tileBytes >> 28);
    {
      const u32 bar = smem_addr(&sBar[buf]);
      const u32 parity = (u32)(kFastStages == 1 ? k : k >> 1) & 1u;
      u32 done = 0;
      while (!done) {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
      }
    }
    fast_run_mode(P, S, tIni, false, 0xffffffffu);
    if (tMin < tIni || S.retryMax > 0) {  // cells without a keypoint at iniThFAST run again at minThFAST (:1119; ComputeKeyPointsOld: at most 3, :1343)
      // the cell flags are final since the run's last barrier (every warp is past phase P3); the answer is the same in every warp
      const int ln = tid & 31;
      const u32 retry = __ballot_sync(0xffffffffu, ln < d.nCells && sCellIni[ln] <= S.retryMax);
      if (retry) {
        __syncthreads();  // the m-map is cleared
        fast_run_mode(P, S, tMin, true, retry);
      }
    }
    __syncthreads();
    if (kFastStages == 1 && tid == 0 && item + (int)gridDim.x < nItems) issue(nslot, load_desc(nstrip), 0);
    item += gridDim.x; slot = nslot; strip = nstrip;
  }
}

int launch_fast(const Params& P, const TmaMaps& M, int nSlots, cudaStream_t st) {
  const int nItems = P.nStrips * nSlots;
  k_fast<<<std::min(nItems, P.fastGrid), kFastThreads, fast_smem_bytes(P), st>>>(P, M, nItems);
  return 1;
}

// Called once per context (after cudaSetDevice): opts k_fast into its dynamic shared memory and sizes the persistent grid.
int configure_fast(Params& P) {
  const size_t smem = fast_smem_bytes(P);
  int dev = 0, sms = 0, perSm = 0, optin = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return ORBX_ERR_CUDA;
  // the attribute is per function and device, not per context: opt in to the device maximum once, contexts differ in what they ask for
  if (cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess) return ORBX_ERR_CUDA;
  optin -= 2048;  // the kernel's static shared memory counts against the same limit
  if ((size_t)optin < smem) return ORBX_ERR_CUDA;
  if (cudaFuncSetAttribute(k_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, optin) != cudaSuccess) return ORBX_ERR_CUDA;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_fast, kFastThreads, smem) != cudaSuccess || perSm < 1) return ORBX_ERR_CUDA;
  P.fastGrid = sms * perSm;
  return ORBX_OK;
}

}  // namespace orbx
