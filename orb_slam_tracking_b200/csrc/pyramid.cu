// Image pyramid: level l = cv::resize(level l-1, INTER_LINEAR) in OpenCV's 8U fixed-point
// arithmetic (ORBextractor::ComputePyramid, Features/ORBextractor.cpp:1660-1713; arithmetic in
// SURVEY.md appendix A1).  The 19-px BORDER_REFLECT_101 frame the reference adds is never read
// by any later stage (FAST stays >= 16 px inside, IC_Angle and BRIEF stay >= 1 px inside, the blur
// reflects the level itself), so levels are stored unpadded and the frame is synthesised only
// when a caller asks for mvImagePyramid (orbx_get_pyramid_level).
//
// HBM-bound: reads P_{l-1}, writes P_l.  One warp per strip of 128 output columns x 16 output rows; a lane
// owns 4 adjacent output columns, so everything that depends on the column only is lane-constant: the 8
// horizontal coefficients, and the position of its <= 6 source bytes, which sit in 2 (sometimes 3) aligned
// words -- one funnel shift pair aligns them and two byte permutes gather the left / right taps of the 4
// pixels.  Walking down the rows, the horizontally interpolated source row sy+1 of output row dy is kept in
// registers: at scale 1.2 it is source row sy of row dy+1 five times out of six.  Coefficient tables are
// computed once on the host with the reference's float/double expressions and kept in L2.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace orbx {

constexpr int kResizeRows = 16, kResizeWarps = 4;
constexpr int kChainMaxSlots = 2;  // batches up to this size build their pyramids in one launch (k_pyramid_chain)

// tab: xofs[dw] | xcoef[dw] (a0 | a1 << 16) | per output row { r0 | r1 << 16, b0 | b1 << 16 }
// source words of the row pass: read-only for the whole launch in k_resize (non-coherent path); in k_pyramid_chain the source level
// was written earlier in the SAME launch by other CTAs, so its loads go to L2 (ld.global.cg), where those stores are
template <bool CHAIN>
__device__ __forceinline__ u32 ld_src(const u32* p) { return CHAIN ? __ldcg(p) : __ldg(p); }
template <bool CHAIN>
__device__ __forceinline__ u32 ld_src(const u8* p) { return CHAIN ? (u32)__ldcg(p) : (u32)__ldg(p); }

// One warp: the strip of `rows` output rows from dy0 and 128 output columns from 128 * bx of level l of `slot`.
// FOLD (the last column block of a level when at most 16 of its lanes have columns inside the level, e.g. 533 = 4 * 128 + 21
// columns: 6 lanes): the warp is cut into 32 / lanesPer groups of lanesPer lanes and group g takes the rows from dyWarp + g * rows.
// The instructions a warp issues do not depend on how many of its lanes are inside the level, so the block costs 1 / groups of
// what it did (a sixth of all strips are such blocks at 640x480, scale 1.2).
template <bool WIDE, bool CHAIN, bool FOLD>
__device__ __forceinline__ void resize_strip(const Params& P, int l, const int* __restrict__ tab, int rows, int bx, int dyWarp, int slot, int lanesPer = 32) {
  const LevelInfo& D = P.lv[l];
  const LevelInfo& S = P.lv[l - 1];
  const int lane = FOLD ? (int)threadIdx.x & (lanesPer - 1) : (int)threadIdx.x;
  const int dx0 = (bx * 32 + lane) * 4;
  if (dyWarp >= D.h) return;  // the whole warp
  const int dy0 = FOLD ? dyWarp + ((int)threadIdx.x / lanesPer) * rows : dyWarp;  // groups beyond the level have no rows below
  const u8* src = level_ptr((const u8*)P.pyr, P, slot, l - 1);
  u8* dst = level_ptr(P.pyr, P, slot, l);
  // bases and pitches the compiler has to keep in registers (it reloaded them from the parameter block for every row)
  int sPitch = S.pitch, dPitch = D.pitch;
  asm volatile("" : "+l"(src), "+l"(dst), "+r"(sPitch), "+r"(dPitch));
  const int* xofs = tab;
  const int* xcoef = tab + D.w;
  const int2* __restrict__ ytab = reinterpret_cast<const int2*>(tab + 2 * D.w);  // the table base is 256-byte aligned
  // lane constants
  int a0[4], a1[4], e0[4], e1[4];
  u32 acoef[4];
  int sx0 = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int dx = min(dx0 + k, D.w - 1);  // columns beyond the level repeat the last one and are not stored
    const int sx = __ldg(xofs + dx), ac = __ldg(xcoef + dx);
    if (k == 0) sx0 = sx;
    a0[k] = (short)(ac & 0xffff); a1[k] = ac >> 16;
    e0[k] = sx - sx0; e1[k] = min(sx + 1, S.w - 1) - sx0;
    acoef[k] = (u32)ac;                                        // a0 | a1 << 16: the two 16-bit operands of dp2a
  }
  // one byte permute per PAIR of pixels: bytes (S[sx_k], S[sx_k + 1], S[sx_k+1], S[sx_k+1 + 1]) out of the 8-byte window; dp2a.lo
  // takes the first two, dp2a.hi the other two
  u32 sel[2];
#pragma unroll
  for (int k = 0; k < 2; ++k)
    sel[k] = (u32)(e0[2 * k] & 7) | ((u32)(e1[2 * k] & 7) << 4) | ((u32)(e0[2 * k + 1] & 7) << 8) | ((u32)(e1[2 * k + 1] & 7) << 12);

  const int A = sx0 & ~3, sh = 8 * (sx0 & 3);
  const bool need3 = (sx0 & 3) + e1[3] >= 8;

  // WIDE: scale factors above 2 spread a lane's taps over more than the 8 bytes the permutes can reach -> byte loads
  // horizontally interpolated row r, pre-shifted: (S[sx] * a0 + S[sx + 1] * a1) >> 4; the two taps are adjacent source
  // bytes, so one byte permute puts them side by side and one dp2a (2 x u16 . 2 x u8) is the whole interpolation
  auto hload = [&](int r, u32 (&w)[3]) {  // the 2 (3) aligned words of source row r that hold this lane's taps
    const u32* rp = reinterpret_cast<const u32*>(src + (u32)(A + r * sPitch));  // base + 32-bit offset
    w[0] = ld_src<CHAIN>(rp); w[1] = ld_src<CHAIN>(rp + 1); w[2] = need3 ? ld_src<CHAIN>(rp + 2) : 0u;
  };
  auto hcalc = [&](const u32 (&w)[3], u32 (&hs)[4]) {
    const u32 q0 = __funnelshift_r(w[0], w[1], sh), q1 = __funnelshift_r(w[1], w[2], sh);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const u32 taps = __byte_perm(q0, q1, sel[k]);
      u32 d0, d1;
      asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d0) : "r"(acoef[2 * k]), "r"(taps), "r"(0u));
      asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d1) : "r"(acoef[2 * k + 1]), "r"(taps), "r"(0u));
      hs[2 * k] = d0 >> 4; hs[2 * k + 1] = d1 >> 4;
    }
  };
  auto hrow = [&](int r, u32 (&hs)[4]) {
    if (WIDE) {
      const u8* rowp = src + r * sPitch + sx0;
#pragma unroll
      for (int k = 0; k < 4; ++k) hs[k] = (ld_src<CHAIN>(rowp + e0[k]) * (u32)a0[k] + ld_src<CHAIN>(rowp + e1[k]) * (u32)a1[k]) >> 4;
      return;
    }
    u32 w[3];
    hload(r, w);
    hcalc(w, hs);
  };

  const int dyEnd = min(dy0 + rows, D.h);
  if (CHAIN && !WIDE && rows == 4) {
    // latency form (k_pyramid_chain): the walk below is a chain of dependent loads per row (row table entry, then source
    // words); here all four table entries and then all eight source rows are requested before anything is used
    int2 yt[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) yt[k] = __ldg(ytab + min(dy0 + k, D.h - 1));
    u32 wa[4][3], wb[4][3];
#pragma unroll
    for (int k = 0; k < 4; ++k) { hload(yt[k].x & 0xffff, wa[k]); hload((int)((u32)yt[k].x >> 16), wb[k]); }
    u8* dq = dst + (size_t)dy0 * dPitch + dx0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (dy0 + k < D.h) {
        u32 h0[4], h1[4], t[2];
        hcalc(wa[k], h0);
        hcalc(wb[k], h1);
        const u32 B0 = (u32)yt[k].y << 16, B1 = (u32)yt[k].y & 0xffff0000u;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const u32 x = __byte_perm(__umulhi(B0, h0[2 * j]), __umulhi(B0, h0[2 * j + 1]), 0x5410);
          const u32 y = __byte_perm(__umulhi(B1, h1[2 * j]), __umulhi(B1, h1[2 * j + 1]), 0x5410);
          t[j] = (x + y + 0x00020002u) >> 2;
        }
        __stcg(reinterpret_cast<u32*>(dq), __byte_perm(t[0], t[1], 0x6420));
      }
      dq += dPitch;
    }
    return;
  }
  {
    // the strip's source rows are consumed one after the other (a dependent chain of loads per lane): pull its
    // 128-byte lines into L2 up front, one line per lane and round
    const int dyWarpEnd = FOLD ? min(dyWarp + (32 / lanesPer) * rows, D.h) : dyEnd;  // all groups' rows
    const int rFirst = __ldg(&ytab[dyWarp].x) & 0xffff, rLast = (int)((u32)__ldg(&ytab[dyWarpEnd - 1].x) >> 16);
    const int xb = __shfl_sync(0xffffffffu, sx0, 0) & ~127;
    const int nLines = (min(__shfl_sync(0xffffffffu, sx0 + e1[3], 31), S.pitch - 1) - xb) / 128 + 1;
    const int total = (rLast - rFirst + 1) * nLines;
    for (int i = threadIdx.x; i < total; i += 32) {
      const int r = i / nLines, ln = i - r * nLines;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(src + (size_t)(rFirst + r) * sPitch + xb + 128 * ln));
    }
  }
  // Output row dy from the horizontally interpolated source rows r0, r1.  At scale 1.2 row r1 of dy is row r0 of dy + 1 five
  // times out of six, so the rows alternate between two register sets (hHave holds row rHave on entry, hNew row r1 on exit;
  // the caller swaps the roles: no register moves).
  int rHave = -1;
  // pitch = width rounded up to 128 and a warp covers 128 columns: every lane's 4-pixel group is inside the row allocation,
  // lanes beyond the level's width write row padding that nothing reads
  u8* dp = dst + (size_t)dy0 * dPitch + dx0;
  auto out_row = [&](int dy, u32 (&hHave)[4], u32 (&hNew)[4]) {
    const int2 yt = __ldg(ytab + dy);
    const int r0 = yt.x & 0xffff, r1 = (int)((u32)yt.x >> 16);
    // (b * h) >> 16 == umulhi(b << 16, h): coefficients are in [0, 2048], h < 2^15
    const u32 B0 = (u32)yt.y << 16, B1 = (u32)yt.y & 0xffff0000u;
    if (r0 != rHave) hrow(r0, hHave);
    hrow(r1, hNew);  // r1 == r0 only where the row loop clips (never when shrinking): the row is simply interpolated again
    // (umulhi(B0, h0) + umulhi(B1, h1) + 2) >> 2 per pixel (<= 255): the products (<= 1020) of two pixels are put side by side in
    // 16-bit halves first, so one add and one shift serve a pixel pair and the final permute picks bytes 0 and 2 of each pair
    u32 s[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const u32 x = __byte_perm(__umulhi(B0, hHave[2 * k]), __umulhi(B0, hHave[2 * k + 1]), 0x5410);
      const u32 y = __byte_perm(__umulhi(B1, hNew[2 * k]), __umulhi(B1, hNew[2 * k + 1]), 0x5410);
      s[k] = (x + y + 0x00020002u) >> 2;
    }
    __stcg(reinterpret_cast<u32*>(dp), __byte_perm(s[0], s[1], 0x6420));
    dp += dPitch;
    rHave = r1;
  };
  u32 hA[4] = {0, 0, 0, 0}, hB[4] = {0, 0, 0, 0};
  for (int dy = dy0; dy < dyEnd; dy += 2) {
    out_row(dy, hA, hB);
    if (dy + 1 >= dyEnd) break;
    out_row(dy + 1, hB, hA);
  }
}

template <bool WIDE>
__global__ void __launch_bounds__(32 * kResizeWarps) k_resize(const __grid_constant__ Params P, int l, const int* __restrict__ tab, int rows) {
  const int warp = blockIdx.y * kResizeWarps + threadIdx.y;
  if (blockIdx.x + 1 == gridDim.x) {
    const int nl = (P.lv[l].w - 128 * (int)blockIdx.x + 3) >> 2;  // lanes of the last column block with columns inside the level (>= 1)
    if (nl <= 16) {
      int lanesPer = 16;
      while ((lanesPer >> 1) >= nl) lanesPer >>= 1;
      resize_strip<WIDE, false, true>(P, l, tab, rows, blockIdx.x, warp * (32 / lanesPer) * rows, blockIdx.z, lanesPer);
      return;
    }
  }
  resize_strip<WIDE, false, false>(P, l, tab, rows, blockIdx.x, warp * rows, blockIdx.z);
}

// The whole pyramid of a few frames in ONE launch (single-frame latency path): seven dependent launches cost about 10 us each
// for a frame, most of it launch latency and the short dependent load chain of a nearly empty GPU.  All CTAs are co-resident (the
// grid is at most one CTA per SM); a level is finished by a grid-wide barrier -- an arrival counter in global memory that the host
// zeroes in front of the launch -- before the next level reads it.
struct ResizeTabs { const int* t[kMaxLevels]; };
__global__ void __launch_bounds__(32 * kResizeWarps) k_pyramid_chain(const __grid_constant__ Params P, const __grid_constant__ ResizeTabs T, int rows,
                                                                     int nSlots, u32* __restrict__ arrive) {
  const int warpsTotal = (int)gridDim.x * kResizeWarps, warp0 = (int)blockIdx.x * kResizeWarps + (int)threadIdx.y;
  for (int l = 1; l < P.nlevels; ++l) {
    const int nbx = (P.lv[l].w + 127) / 128, nwy = (P.lv[l].h + rows - 1) / rows, perSlot = nbx * nwy;
    const bool wide = P.lv[l - 1].w > 2 * P.lv[l].w;  // horizontal scale above 2 (by however little: 309 -> 154 columns)
    for (int it = warp0; it < perSlot * nSlots; it += warpsTotal) {
      const int slot = it / perSlot, r = it - slot * perSlot, wy = r / nbx, bx = r - wy * nbx;
      if (wide) resize_strip<true, true, false>(P, l, T.t[l], rows, bx, wy * rows, slot);
      else resize_strip<false, true, false>(P, l, T.t[l], rows, bx, wy * rows, slot);
    }
    if (l + 1 < P.nlevels) {  // grid barrier: every store of level l is in L2 before any CTA starts level l + 1
      __syncthreads();
      if (threadIdx.x == 0 && threadIdx.y == 0) {
        __threadfence();
        atomicAdd(arrive, 1u);
        const u32 want = (u32)l * gridDim.x;
        u32 seen;
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(arrive) : "memory");
        } while (seen < want);
      }
      __syncthreads();
    }
  }
}

// Level 0 of each slot from a dense [n][H][W] staging buffer (one DMA from the host for a whole chunk of
// frames instead of one strided copy per frame): 16 bytes per thread where the geometry allows.
__global__ void __launch_bounds__(256) k_ingest(const __grid_constant__ Params P, const u8* __restrict__ dense) {
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[0];
  const u8* __restrict__ src = dense + (size_t)slot * P.W * P.H;
  u8* dst = level_ptr(P.pyr, P, slot, 0);
  if ((P.W & 15) == 0) {
    const int wv = P.W >> 4, total = wv * P.H;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
      const int y = i / wv, x = i - y * wv;
      reinterpret_cast<uint4*>(dst + (size_t)y * L.pitch)[x] = __ldg(reinterpret_cast<const uint4*>(src + (size_t)y * P.W) + x);
    }
  } else {
    const int total = P.W * P.H;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
      const int y = i / P.W, x = i - y * P.W;
      dst[(size_t)y * L.pitch + x] = __ldg(src + i);
    }
  }
}

// Converter::toGray (Utils/Converter.cpp:5-19) / cv::cvtColor(BGR2GRAY | RGB2GRAY) (demo/demo_initialization.cpp:67-68)
// fused into the ingest: level 0 = (3735 B + 19235 G + 9798 R + 16384) >> 15 (OpenCV's 8U fixed point)
// from a dense interleaved [n][H][W][3] staging buffer.  One thread per 4 output pixels (three 32-bit loads).
__global__ void __launch_bounds__(256) k_ingest_color(const __grid_constant__ Params P, const u8* __restrict__ dense, int rgb) {
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[0];
  const u8* __restrict__ src = dense + (size_t)slot * P.W * P.H * 3;
  u8* dst = level_ptr(P.pyr, P, slot, 0);
  const int c0 = rgb ? 9798 : 3735, c2 = rgb ? 3735 : 9798;  // weight of channel 0 / channel 2
  const int wq = (P.W + 3) >> 2, total = wq * P.H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int y = i / wq, x = (i - y * wq) * 4;
    const u8* s = src + ((size_t)y * P.W + x) * 3;
    u32 out = 0;
    if (x + 3 < P.W && ((reinterpret_cast<size_t>(s) & 3) == 0)) {
      const u32 w0 = __ldg(reinterpret_cast<const u32*>(s)), w1 = __ldg(reinterpret_cast<const u32*>(s) + 1),
                w2 = __ldg(reinterpret_cast<const u32*>(s) + 2);
      // bytes: w0 = b0 g0 r0 b1 | w1 = g1 r1 b2 g2 | w2 = r2 b3 g3 r3
      const int g0 = (c0 * (int)(w0 & 255) + 19235 * (int)((w0 >> 8) & 255) + c2 * (int)((w0 >> 16) & 255) + 16384) >> 15;
      const int g1 = (c0 * (int)(w0 >> 24) + 19235 * (int)(w1 & 255) + c2 * (int)((w1 >> 8) & 255) + 16384) >> 15;
      const int g2 = (c0 * (int)((w1 >> 16) & 255) + 19235 * (int)(w1 >> 24) + c2 * (int)(w2 & 255) + 16384) >> 15;
      const int g3 = (c0 * (int)((w2 >> 8) & 255) + 19235 * (int)((w2 >> 16) & 255) + c2 * (int)(w2 >> 24) + 16384) >> 15;
      out = (u32)g0 | ((u32)g1 << 8) | ((u32)g2 << 16) | ((u32)g3 << 24);
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (x + k < P.W) {
          const int g = (c0 * (int)__ldg(s + 3 * k) + 19235 * (int)__ldg(s + 3 * k + 1) + c2 * (int)__ldg(s + 3 * k + 2) + 16384) >> 15;
          out |= (u32)g << (8 * k);
        }
    }
    *reinterpret_cast<u32*>(dst + (size_t)y * L.pitch + x) = out;  // pitch is a multiple of 128: the word is inside the row
  }
}

int launch_ingest_color(const Params& P, const u8* dense, int rgb, int nSlots, cudaStream_t st) {
  const int work = ((P.W + 3) >> 2) * P.H;
  dim3 grid(std::min((work + 255) / 256, 128), nSlots);
  k_ingest_color<<<grid, 256, 0, st>>>(P, dense, rgb);
  return 1;
}

int launch_ingest(const Params& P, const u8* dense, int nSlots, cudaStream_t st) {
  const int work = (P.W & 15) == 0 ? (P.W >> 4) * P.H : P.W * P.H;
  dim3 grid(std::min((work + 255) / 256, 64), nSlots);
  k_ingest<<<grid, 256, 0, st>>>(P, dense);
  return 1;
}

int launch_pyramid(const Params& P, const int* const* dTabs, int nSlots, cudaStream_t st) {
  int n = 0;
  if (nSlots <= kChainMaxSlots && P.optPyramidChain && P.gridBar) {
    // few frames: one launch for all levels (k_pyramid_chain).  4 rows per warp keep the dependent load chain short; the grid is
    // what level 1 needs, capped at one CTA per SM so that every CTA is resident (the barrier spins)
    const int rows = P.optResizeRows > 0 ? P.optResizeRows : 4;
    const int items = ((P.lv[1].w + 127) / 128) * ((P.lv[1].h + rows - 1) / rows) * nSlots;
    const int grid = std::max(1, std::min(P.smCount, (items + kResizeWarps - 1) / kResizeWarps));
    ResizeTabs T;
    for (int l = 0; l < kMaxLevels; ++l) T.t[l] = l < P.nlevels ? dTabs[l] : nullptr;
    cudaMemsetAsync(P.gridBar, 0, sizeof(u32), st);
    k_pyramid_chain<<<grid, dim3(32, kResizeWarps), 0, st>>>(P, T, rows, nSlots, P.gridBar);
    return 1;
  }
  for (int l = 1; l < P.nlevels; ++l) {
    dim3 block(32, kResizeWarps);
    // a warp walks its rows one after the other (a dependent chain of loads): with whole batches the SMs are full anyway and
    // 16 rows per warp amortise the lane constants best; a single frame wants the chain short (its latency is the kernel's)
    const int rows = P.optResizeRows > 0 ? P.optResizeRows : nSlots >= 64 ? kResizeRows : nSlots >= 8 ? 8 : 4;
    dim3 grid((P.lv[l].w + 127) / 128, (P.lv[l].h + rows * kResizeWarps - 1) / (rows * kResizeWarps), nSlots);
    // taps of 4 adjacent output columns span 3 * scale + 2 source bytes
    // the permutes reach 8 bytes = the taps of 4 columns up to a horizontal scale of exactly 2; anything above (309 -> 154 columns is
    // 2.006) takes the byte-load form
    if (P.lv[l - 1].w > 2 * P.lv[l].w) k_resize<true><<<grid, block, 0, st>>>(P, l, dTabs[l], rows);
    else k_resize<false><<<grid, block, 0, st>>>(P, l, dTabs[l], rows);
    ++n;
  }
  return n;
}

}  // namespace orbx
