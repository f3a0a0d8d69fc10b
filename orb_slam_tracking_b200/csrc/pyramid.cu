// Image pyramid: level l = cv::resize(level l-1, INTER_LINEAR) in OpenCV's 8U fixed-point
// arithmetic (ORBextractor::ComputePyramid, Features/ORBextractor.cpp:1660-1713; arithmetic in
// SURVEY.md appendix A1).  The 19-px BORDER_REFLECT_101 frame the reference adds is never read
// by any later stage (FAST stays >= 16 px inside, IC_Angle and BRIEF stay >= 1 px inside, the blur
// reflects the level itself), so levels are stored unpadded and the frame is synthesised only
// when a caller asks for mvImagePyramid (orbx_get_pyramid_level).
//
// HBM-bound: reads P_{l-1}, writes P_l.  Each thread produces 4 adjacent pixels and stores one
// 32-bit word; the two source rows are read through the read-only path (a warp touches two
// 154-byte spans per row).  Coefficient tables are computed once on the host with the
// reference's float/double expressions and kept in L2.
#include <algorithm>

#include "common.cuh"

namespace orbx {

// tab: xofs[dw] | xcoef[dw] (a0 | a1 << 16) | yofs[dh] | ycoef[dh] (b0 | b1 << 16)
__global__ void __launch_bounds__(256) k_resize(const __grid_constant__ Params P, int l, const int* __restrict__ tab) {
  const LevelInfo& D = P.lv[l];
  const LevelInfo& S = P.lv[l - 1];
  const int slot = blockIdx.z;
  const int dy = blockIdx.y * blockDim.y + threadIdx.y;
  const int dx0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (dy >= D.h || dx0 >= D.w) return;
  const u8* __restrict__ src = level_ptr((const u8*)P.pyr, P, slot, l - 1);
  u8* dst = level_ptr(P.pyr, P, slot, l);
  const int* xofs = tab;
  const int* xcoef = tab + D.w;
  const int* yofs = tab + 2 * D.w;
  const int* ycoef = yofs + D.h;
  int sy = __ldg(yofs + dy);
  const int bc = __ldg(ycoef + dy);
  const int b0 = (short)(bc & 0xffff), b1 = bc >> 16;
  const int sy0 = min(max(sy, 0), S.h - 1), sy1 = min(max(sy + 1, 0), S.h - 1);
  const u8* r0 = src + (size_t)sy0 * S.pitch;
  const u8* r1 = src + (size_t)sy1 * S.pitch;
  u32 packed = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int dx = dx0 + k;
    if (dx < D.w) {
      const int sx = __ldg(xofs + dx);
      const int ac = __ldg(xcoef + dx);
      const int a0 = (short)(ac & 0xffff), a1 = ac >> 16;
      const int sx1 = min(sx + 1, S.w - 1);
      const int h0 = __ldg(r0 + sx) * a0 + __ldg(r0 + sx1) * a1;
      const int h1 = __ldg(r1 + sx) * a0 + __ldg(r1 + sx1) * a1;
      const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
      packed |= (u32)(v & 0xff) << (8 * k);
    }
  }
  // pitch is a multiple of 128, so the 4-pixel group is always inside the row allocation
  *reinterpret_cast<u32*>(dst + (size_t)dy * D.pitch + dx0) = packed;
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

int launch_ingest(const Params& P, const u8* dense, int nSlots, cudaStream_t st) {
  const int work = (P.W & 15) == 0 ? (P.W >> 4) * P.H : P.W * P.H;
  dim3 grid(std::min((work + 255) / 256, 64), nSlots);
  k_ingest<<<grid, 256, 0, st>>>(P, dense);
  return 1;
}

int launch_pyramid(const Params& P, const int* const* dTabs, int nSlots, cudaStream_t st) {
  int n = 0;
  for (int l = 1; l < P.nlevels; ++l) {
    dim3 block(32, 8);
    dim3 grid((P.lv[l].w + 127) / 128, (P.lv[l].h + 7) / 8, nSlots);
    k_resize<<<grid, block, 0, st>>>(P, l, dTabs[l]);
    ++n;
  }
  return n;
}

}  // namespace orbx
