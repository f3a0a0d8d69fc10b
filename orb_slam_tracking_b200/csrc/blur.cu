// cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) in OpenCV's 8U fixed-point arithmetic
// (call site Features/ORBextractor.cpp:1598-1606; arithmetic in SURVEY.md appendix A5):
// separable integer kernel {18,34,48,56,48,34,18}/256, horizontal sums in 8.8 (<= 65280, u16),
// vertical sums in 8.16, out = (v + 32768) >> 16.  The reference blurs a clone of the level, so the
// border reflects the level itself.
//
// All levels of all slots in ONE launch (host-built strip table).  HBM-bound by design: reads P_l
// (+ halo out of L2), writes P_l.  One WARP per strip of 120 columns x 64 rows, no shared memory:
// lane = one aligned 32-bit word (4 columns) of the 128 staged columns x0-4 .. x0+123; lanes 1..30 produce
// output, lanes 0 and 31 only carry the halo.  Walking down the rows, a lane loads its word (rows are
// fetched kRB = 10 at a time so that the loads of a block are in flight together), takes the neighbouring words
// from the adjacent lanes by shuffle, runs the row pass with funnel shifts and dp4a, and keeps the last 7
// row-pass results of its 4 columns in registers for the column pass (rows paired in 16-bit halves:
// 3 dp2a + 1 multiply-add per pixel; the window of five row pairs is a ring whose position is a compile-time constant
// because the row block is a multiple of five rows: no register moves).  BORDER_REFLECT_101 in x costs nothing in the loop: a word that touches or
// lies beyond the level's border is a lane-constant byte permutation of two aligned words inside the level.
#include "common.cuh"

namespace orbx {

constexpr int kTW = kBlurTW, kTH = kBlurTH;
#ifndef ORBX_BLUR_WARPS
#define ORBX_BLUR_WARPS 4
#endif
#ifndef ORBX_BLUR_RB
#define ORBX_BLUR_RB 10
#endif
constexpr int kBlurWarps = ORBX_BLUR_WARPS;
constexpr int kRB = ORBX_BLUR_RB;  // rows per block of loads: a multiple of the 5-deep pair ring, and (kTH + 6) % kRB == 0
static_assert((kTH + 6) % kRB == 0 && kRB % 5 == 0, "row block");

__global__ void __launch_bounds__(kBlurWarps * 32) k_blur(const __grid_constant__ Params P) {
  const int lane = threadIdx.x & 31;
  const int ti = blockIdx.x * kBlurWarps + (threadIdx.x >> 5);
  if (ti >= P.nBlurTiles) return;
  const BlurTile t = P.blurTiles[ti];
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[t.level];
  const u8* __restrict__ src = level_ptr((const u8*)P.pyr, P, slot, t.level);
  u8* dst = level_ptr(P.blur, P, slot, t.level);
  const int xw = t.x0 - 4 + 4 * lane;  // my 4 columns
  // my word = bytes at the reflected columns c_k, all inside two aligned words starting at column A
  int A, two;
  u32 S;
  {
    int c[4], lo = 1 << 30, hi = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int v = xw + k;
      v = v < 0 ? -v : v;
      v = v >= L.w ? 2 * (L.w - 1) - v : v;
      v = min(max(v, 0), L.w - 1);  // lanes far beyond the level: any valid address, never used
      c[k] = v; lo = min(lo, v); hi = max(hi, v);
    }
    A = lo & ~3;
    S = (u32)(c[0] - A) | ((u32)(c[1] - A) << 4) | ((u32)(c[2] - A) << 8) | ((u32)(c[3] - A) << 12);
    two = hi - A >= 4;
  }
  const bool emit = lane >= 1 && lane <= 30 && xw < L.w;
  constexpr u32 kA = 18u | (34u << 8) | (48u << 16) | (56u << 24);  // taps 0..3
  constexpr u32 kB = 48u | (34u << 8) | (18u << 16);                // taps 4..6 (+ one ignored byte)
  // column pass window: the row-pass results (<= 65280, 16 bits) of consecutive rows are kept PAIRED per pixel,
  // pr[q][j] = h(row q) | h(row q + 1) << 16, so that dp2a (2 x u16 . 2 x u8) applies two taps per instruction
  constexpr u32 kT01 = 18u | (34u << 8), kT23 = 48u | (56u << 8), kT45 = 48u | (34u << 8);
  u32 pr[5][4], hl[4];
#pragma unroll
  for (int k = 0; k < 5; ++k) { pr[k][0] = pr[k][1] = pr[k][2] = pr[k][3] = 0; }
  hl[0] = hl[1] = hl[2] = hl[3] = 0;
  // The lane's column of the source and of the destination as pointers the compiler has to keep (and walk down the rows):
  // left alone it rebuilds every row address from the parameter block (a constant-bank load and two 64-bit additions per row).
  const u8* colp = src + A;
  u8* dcol = dst + xw;
  int pitch = L.pitch;
  asm volatile("" : "+l"(colp), "+l"(dcol), "+r"(pitch));
  for (int rb = 0; rb < kTH + 6; rb += kRB) {
    if (t.y0 + rb - 6 >= L.h) break;  // nothing left to emit in this strip
    u32 v0[kRB], v1[kRB];
    const int yb = t.y0 + rb - 3;
    if (yb >= 0 && yb + kRB <= L.h) {  // all rows of the block inside the level: one pointer walks down
      const u8* rp = colp + yb * pitch;
#pragma unroll
      for (int k = 0; k < kRB; ++k, rp += pitch) {
        v0[k] = __ldg(reinterpret_cast<const u32*>(rp));
        v1[k] = two ? __ldg(reinterpret_cast<const u32*>(rp) + 1) : 0u;
      }
    } else {
#pragma unroll
      for (int k = 0; k < kRB; ++k) {
        int y = yb + k;
        y = y < 0 ? -y : y;
        y = y >= L.h ? 2 * (L.h - 1) - y : y;
        y = max(y, 0);  // rows beyond the halo of the last strip are never used
        const u32* rp = reinterpret_cast<const u32*>(colp + y * pitch);
        v0[k] = __ldg(rp);
        v1[k] = two ? __ldg(rp + 1) : 0u;
      }
    }
    const int nEmit = emit ? L.h - (t.y0 + rb - 6) : 0;  // rows k < nEmit of the block are inside the level
    u8* dp = dcol + (t.y0 + rb - 6) * pitch;
#pragma unroll
    for (int k = 0; k < kRB; ++k) {
      const u32 w1 = __byte_perm(v0[k], v1[k], S);
      const u32 w0 = __shfl_up_sync(0xffffffffu, w1, 1);
      const u32 w2 = __shfl_down_sync(0xffffffffu, w1, 1);
      // row pass: pixel j uses bytes (1+j .. 4+j) of (w0,w1) and (1+j .. 4+j) of (w1,w2)
      const u32 n0 = __dp4a(__funnelshift_r(w0, w1, 8), kA, __dp4a(__funnelshift_r(w1, w2, 8), kB, 0u));
      const u32 n1 = __dp4a(__funnelshift_r(w0, w1, 16), kA, __dp4a(__funnelshift_r(w1, w2, 16), kB, 0u));
      const u32 n2 = __dp4a(__funnelshift_r(w0, w1, 24), kA, __dp4a(__funnelshift_r(w1, w2, 24), kB, 0u));
      const u32 n3 = __dp4a(w1, kA, __dp4a(w2, kB, 0u));
      const u32 nn[4] = {n0, n1, n2, n3};
      u32 px[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // rows 0..6 of the window = ring entries 0 (rows 0,1), 2 (2,3), 4 (4,5) and the new row; entry i is pr[(i + k) % 5]
        u32 v;
        asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(v) : "r"(pr[k % 5][j]), "r"(kT01), "r"(18u * nn[j] + 32768u));
        asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(v) : "r"(pr[(k + 2) % 5][j]), "r"(kT23), "r"(v));
        asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(v) : "r"(pr[(k + 4) % 5][j]), "r"(kT45), "r"(v));
        px[j] = v;  // < 2^24: the output pixel is byte 2
        pr[k % 5][j] = __byte_perm(hl[j], nn[j], 0x5410);  // (previous row, new row) replaces the oldest pair
        hl[j] = nn[j];
      }
      if (k < nEmit && (k >= 6 || rb > 0)) {
        __stcg(reinterpret_cast<u32*>(dp), __byte_perm(__byte_perm(px[0], px[1], 0x0062), __byte_perm(px[2], px[3], 0x0062), 0x5410));
      }
      dp += pitch;
    }
  }
}

int launch_blur(const Params& P, int nSlots, cudaStream_t st) {
  dim3 grid((P.nBlurTiles + kBlurWarps - 1) / kBlurWarps, nSlots);
  k_blur<<<grid, kBlurWarps * 32, 0, st>>>(P);
  return 1;
}

}  // namespace orbx
