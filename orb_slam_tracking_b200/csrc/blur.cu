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
// fetched 7 at a time so that the loads of a block are in flight together), takes the neighbouring words
// from the adjacent lanes by shuffle, runs the row pass with funnel shifts and dp4a, and keeps the last 7
// row-pass results of its 4 columns in registers for the column pass (symmetric taps: 3 adds + 4
// multiply-adds per pixel).  BORDER_REFLECT_101 in x costs nothing in the loop: a word that touches or
// lies beyond the level's border is a lane-constant byte permutation of two aligned words inside the level.
#include "common.cuh"

namespace orbx {

constexpr int kTW = 120, kTH = 64;
constexpr int kBlurWarps = 4;

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
  u32 h[7][4];
#pragma unroll
  for (int k = 0; k < 7; ++k) { h[k][0] = h[k][1] = h[k][2] = h[k][3] = 0; }
  const u8* __restrict__ colp = src + A;
  for (int rb = 0; rb < kTH + 6; rb += 7) {
    if (t.y0 + rb - 6 >= L.h) break;  // nothing left to emit in this strip
    u32 v0[7], v1[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      int y = t.y0 + rb + k - 3;
      y = y < 0 ? -y : y;
      y = y >= L.h ? 2 * (L.h - 1) - y : y;
      y = max(y, 0);  // rows beyond the halo of the last strip are never used
      const u32* rp = reinterpret_cast<const u32*>(colp + y * L.pitch);
      v0[k] = __ldg(rp);
      v1[k] = two ? __ldg(rp + 1) : 0u;
    }
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const u32 w1 = __byte_perm(v0[k], v1[k], S);
      const u32 w0 = __shfl_up_sync(0xffffffffu, w1, 1);
      const u32 w2 = __shfl_down_sync(0xffffffffu, w1, 1);
      // row pass: pixel j uses bytes (1+j .. 4+j) of (w0,w1) and (1+j .. 4+j) of (w1,w2)
      const u32 n0 = __dp4a(__funnelshift_r(w0, w1, 8), kA, __dp4a(__funnelshift_r(w1, w2, 8), kB, 0u));
      const u32 n1 = __dp4a(__funnelshift_r(w0, w1, 16), kA, __dp4a(__funnelshift_r(w1, w2, 16), kB, 0u));
      const u32 n2 = __dp4a(__funnelshift_r(w0, w1, 24), kA, __dp4a(__funnelshift_r(w1, w2, 24), kB, 0u));
      const u32 n3 = __dp4a(w1, kA, __dp4a(w2, kB, 0u));
#pragma unroll
      for (int q = 0; q < 6; ++q) { h[q][0] = h[q + 1][0]; h[q][1] = h[q + 1][1]; h[q][2] = h[q + 1][2]; h[q][3] = h[q + 1][3]; }
      h[6][0] = n0; h[6][1] = n1; h[6][2] = n2; h[6][3] = n3;
      const int yo = t.y0 + rb + k - 6;
      if (emit && rb + k >= 6 && yo < L.h) {
        u32 px[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const u32 v = 18u * (h[0][j] + h[6][j]) + 34u * (h[1][j] + h[5][j]) + 48u * (h[2][j] + h[4][j]) + 56u * h[3][j] + 32768u;
          px[j] = v >> 16;
        }
        *reinterpret_cast<u32*>(dst + yo * L.pitch + xw) = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
      }
    }
  }
}

int launch_blur(const Params& P, int nSlots, cudaStream_t st) {
  dim3 grid((P.nBlurTiles + kBlurWarps - 1) / kBlurWarps, nSlots);
  k_blur<<<grid, kBlurWarps * 32, 0, st>>>(P);
  return 1;
}

}  // namespace orbx
