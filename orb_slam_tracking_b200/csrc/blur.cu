// cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) in OpenCV's 8U fixed-point arithmetic
// (call site Features/ORBextractor.cpp:1598-1606; arithmetic in SURVEY.md appendix A5):
// separable integer kernel {18,34,48,56,48,34,18}/256, horizontal sums in 8.8 (<= 65280, u16),
// vertical sums in 8.16, out = (v + 32768) >> 16.  The reference blurs a clone of the level, so the
// border reflects the level itself.
//
// All levels of all slots in ONE launch: the host builds a tile table (level, x0, y0).  HBM-bound:
// reads P_l (+ halo from L2), writes P_l.  A CTA stages (TH+6) x (TW+6) source bytes in shared
// memory (reflect indexing resolved while staging), runs the row pass into a u16 tile and the
// column pass out of it; each thread emits 4 adjacent pixels as one 32-bit store.
#include "common.cuh"

namespace orbx {

constexpr int kTW = 128, kTH = 32, kSW = kTW + 8;  // staged row = 3 halo + 128 + 3 halo (+2 pad)

__global__ void __launch_bounds__(256) k_blur(const __grid_constant__ Params P) {
  __shared__ __align__(16) u8 sSrc[(kTH + 6) * kSW];
  __shared__ __align__(16) unsigned short sH[(kTH + 6) * kTW];
  const BlurTile t = P.blurTiles[blockIdx.x];
  const int slot = blockIdx.y;
  const LevelInfo& L = P.lv[t.level];
  const u8* __restrict__ src = level_ptr((const u8*)P.pyr, P, slot, t.level);
  u8* dst = level_ptr(P.blur, P, slot, t.level);
  const int tid = threadIdx.x;
  // stage source with reflected borders
  for (int i = tid; i < (kTH + 6) * (kTW + 6); i += 256) {
    const int r = i / (kTW + 6), c = i - r * (kTW + 6);
    int y = t.y0 + r - 3;
    int x = t.x0 + c - 3;
    u8 v = 0;
    if (x < L.w + 3 && y < L.h + 3) {  // beyond the level (+halo) nothing is ever used
      x = reflect101(x, L.w);
      y = reflect101(y, L.h);
      v = __ldg(src + (size_t)y * L.pitch + x);
    }
    sSrc[r * kSW + c] = v;
  }
  __syncthreads();
  // row pass
  for (int i = tid; i < (kTH + 6) * kTW; i += 256) {
    const int r = i / kTW, c = i - r * kTW;
    const u8* p = sSrc + r * kSW + c;
    const u32 s = 18u * (p[0] + p[6]) + 34u * (p[1] + p[5]) + 48u * (p[2] + p[4]) + 56u * p[3];
    sH[i] = (unsigned short)s;
  }
  __syncthreads();
  // column pass, 4 pixels per thread
  for (int i = tid; i < kTH * (kTW / 4); i += 256) {
    const int r = i / (kTW / 4), c = (i - r * (kTW / 4)) * 4;
    const int y = t.y0 + r, x = t.x0 + c;
    if (y >= L.h || x >= L.w) continue;
    u32 packed = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned short* q = sH + r * kTW + c + k;
      const u32 v = 18u * (q[0] + q[6 * kTW]) + 34u * (q[kTW] + q[5 * kTW]) + 48u * (q[2 * kTW] + q[4 * kTW]) +
                    56u * q[3 * kTW];
      packed |= ((v + 32768u) >> 16) << (8 * k);
    }
    *reinterpret_cast<u32*>(dst + (size_t)y * L.pitch + x) = packed;
  }
}

int launch_blur(const Params& P, int nSlots, cudaStream_t st) {
  dim3 grid(P.nBlurTiles, nSlots);
  k_blur<<<grid, 256, 0, st>>>(P);
  return 1;
}

}  // namespace orbx
