// cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) in OpenCV's 8U fixed-point arithmetic
// (call site Features/ORBextractor.cpp:1598-1606; arithmetic in SURVEY.md appendix A5):
// separable integer kernel {18,34,48,56,48,34,18}/256, horizontal sums in 8.8 (<= 65280, u16),
// vertical sums in 8.16, out = (v + 32768) >> 16.  The reference blurs a clone of the level, so the
// border reflects the level itself.
//
// All levels of all slots in ONE launch (host-built tile table).  HBM-bound by design: reads P_l
// (+ halo out of L2), writes P_l.  A CTA stages (32+6) rows x (128+8) bytes as aligned 32-bit words
// (byte-wise reflection only in words that straddle the level's border), runs the row pass with
// two funnel-shifts + two dp4a per pixel into a packed u16 tile, and the column pass with dp2a on
// vertically paired u16 values; every thread emits 4 adjacent pixels per row as one 32-bit store.
#include "common.cuh"

namespace orbx {

constexpr int kTW = 128, kTH = 32;
constexpr int kSrcWords = kTW / 4 + 2;  // words per staged row: columns x0-4 .. x0+131
constexpr int kRows = kTH + 6;

__device__ __forceinline__ u32 load_word_reflect(const u8* __restrict__ src, const LevelInfo& L, int y, int xw) {
  // 4 bytes at columns xw..xw+3 of row y (already reflected), BORDER_REFLECT_101 in x
  const u8* row = src + (size_t)y * L.pitch;
  if (xw >= 0 && xw + 3 < L.w) return __ldg(reinterpret_cast<const u32*>(row + xw));
  u32 v = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int x = xw + k;
    x = x < 0 ? -x : x;
    x = x >= L.w ? 2 * (L.w - 1) - x : x;
    x = max(x, 0);  // columns beyond the halo of the last tile are never used
    v |= (u32)__ldg(row + x) << (8 * k);
  }
  return v;
}

__global__ void __launch_bounds__(256) k_blur(const __grid_constant__ Params P) {
  __shared__ u32 sSrc[kRows * kSrcWords];
  __shared__ uint2 sH[kRows * (kTW / 4)];  // 4 x u16 per entry
  const BlurTile t = P.blurTiles[blockIdx.x];
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[t.level];
  const u8* __restrict__ src = level_ptr((const u8*)P.pyr, P, slot, t.level);
  u8* dst = level_ptr(P.blur, P, slot, t.level);
  const int tid = threadIdx.x;

  for (int i = tid; i < kRows * kSrcWords; i += 256) {
    const int r = i / kSrcWords, k = i - r * kSrcWords;
    int y = t.y0 + r - 3;
    y = y < 0 ? -y : y;
    y = y >= L.h ? 2 * (L.h - 1) - y : y;
    y = max(y, 0);  // rows beyond the halo of the last tile are never used
    sSrc[i] = load_word_reflect(src, L, y, t.x0 - 4 + 4 * k);
  }
  __syncthreads();

  // row pass: h[x] = sum_i k[i] * p[x + i - 3]; pixel j of group c sits at staged byte 4 + 4c + j
  constexpr u32 kA = 18u | (34u << 8) | (48u << 16) | (56u << 24);  // taps 0..3
  constexpr u32 kB = 48u | (34u << 8) | (18u << 16);                // taps 4..6 (+ one ignored byte)
  for (int i = tid; i < kRows * (kTW / 4); i += 256) {
    const int r = i >> 5, c = i & 31;
    const u32* w = sSrc + r * kSrcWords + c;
    const u32 w0 = w[0], w1 = w[1], w2 = w[2];
    // pixel j: bytes (1+j .. 4+j) of (w0,w1) and bytes (1+j .. 4+j) of (w1,w2)
    const u32 h0 = __dp4a(__funnelshift_r(w0, w1, 8), kA, __dp4a(__funnelshift_r(w1, w2, 8), kB, 0u));
    const u32 h1 = __dp4a(__funnelshift_r(w0, w1, 16), kA, __dp4a(__funnelshift_r(w1, w2, 16), kB, 0u));
    const u32 h2 = __dp4a(__funnelshift_r(w0, w1, 24), kA, __dp4a(__funnelshift_r(w1, w2, 24), kB, 0u));
    const u32 h3 = __dp4a(w1, kA, __dp4a(w2, kB, 0u));
    sH[i] = make_uint2(h0 | (h1 << 16), h2 | (h3 << 16));
  }
  __syncthreads();

  // column pass: thread = 4 columns x 4 rows; vertical pairs (h[r], h[r+1]) feed dp2a
  {
    const int c = tid & 31, rg = tid >> 5;  // rows 4*rg .. 4*rg+3 of the tile
    const int x = t.x0 + 4 * c;
    if (x < L.w) {
      uint2 h[10];
#pragma unroll
      for (int k = 0; k < 10; ++k) h[k] = sH[(4 * rg + k) * (kTW / 4) + c];
      constexpr u32 k01 = 18u | (34u << 8), k23 = 48u | (56u << 8), k45 = 48u | (34u << 8), k6_ = 18u, k_6 = 18u << 8;
#pragma unroll
      for (int o = 0; o < 4; ++o) {
        const int y = t.y0 + 4 * rg + o;
        if (y < L.h) {
          u32 px[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            // pair(k) = (h[k] pixel j, h[k+1] pixel j)
            const u32 sel = (j & 1) ? 0x7632u : 0x5410u;
#define ORBX_PAIR(k) __byte_perm((j < 2) ? h[k].x : h[k].y, (j < 2) ? h[(k) + 1].x : h[(k) + 1].y, sel)
            u32 v = __dp2a_lo(ORBX_PAIR(o), k01, 32768u);
            v = __dp2a_lo(ORBX_PAIR(o + 2), k23, v);
            v = __dp2a_lo(ORBX_PAIR(o + 4), k45, v);
            if (o < 3) v = __dp2a_lo(ORBX_PAIR(o + 6), k6_, v);   // (h[o+6], h[o+7]) * (18, 0)
            else v = __dp2a_lo(ORBX_PAIR(o + 5), k_6, v);          // (h[o+5], h[o+6]) * (0, 18)
#undef ORBX_PAIR
            px[j] = v >> 16;
          }
          *reinterpret_cast<u32*>(dst + (size_t)y * L.pitch + x) = px[0] | (px[1] << 8) | (px[2] << 16) | (px[3] << 24);
        }
      }
    }
  }
}

int launch_blur(const Params& P, int nSlots, cudaStream_t st) {
  dim3 grid(P.nBlurTiles, nSlots);
  k_blur<<<grid, 256, 0, st>>>(P);
  return 1;
}

}  // namespace orbx
