// Frame constructor glue on the device (SlamTypes/Frame.cpp:40-99, 136-161):
//   UndistortKeyPoints -> cv::undistortPoints(pts, K, D, noArray(), K)   (arithmetic: SURVEY.md A6)
//   PosInGrid + mGrid[64][48] assignment, stored as CSR (cell = ix * 48 + iy, indices ascending,
//   which is the push_back order of the reference's loop over i).
// One CTA per slot.  Double arithmetic with separate multiply/add roundings (-fmad=false).
#include "common.cuh"

namespace orbx {

__device__ __forceinline__ void undistort_point(const Params& P, float u, float v, float* ou, float* ov) {
  const double fx = P.K4[0], fy = P.K4[1], cx = P.K4[2], cy = P.K4[3];
  const double k1 = P.D4[0], k2 = P.D4[1], p1 = P.D4[2], p2 = P.D4[3];
  const double ifx = 1. / fx, ify = 1. / fy;
  double x = __dmul_rn((double)u - cx, ifx), y = __dmul_rn((double)v - cy, ify);
  const double x0 = x, y0 = y;
#pragma unroll 1
  for (int j = 0; j < 5; ++j) {
    const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
    // 1 / (1 + ((k3*r2 + k2)*r2 + k1)*r2), k3 = 0
    const double poly = __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(0.0, r2), k2), r2), k1), r2);
    const double icdist = 1. / __dadd_rn(1.0, poly);
    if (icdist < 0) {  // OpenCV 4.x: a negative radial factor (a point far outside the model's range) ends the iteration at the start value
      x = x0; y = y0;
      break;
    }
    const double dx = __dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, p1), x), y),
                                __dmul_rn(p2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
    const double dy = __dadd_rn(__dmul_rn(p1, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                __dmul_rn(__dmul_rn(__dmul_rn(2.0, p2), x), y));
    x = __dmul_rn(x0 - dx, icdist);
    y = __dmul_rn(y0 - dy, icdist);
  }
  *ou = (float)__dadd_rn(__dmul_rn(x, fx), cx);
  *ov = (float)__dadd_rn(__dmul_rn(y, fy), cy);
}

// CTA size GT: 256 threads for whole batches (several CTAs per SM), 1024 for a few frames -- a keypoint per thread for the usual
// 1000, because the undistortion is a long dependent chain in double and a single frame's kernel time is its latency

template <int GT>
__global__ void __launch_bounds__(GT) k_glue(const __grid_constant__ Params P) {
  constexpr int kCellsPerThread = kGridCells / GT;
  static_assert(kGridCells % GT == 0, "cells per thread");
  __shared__ int sCnt[kGridCells];
  __shared__ int sPart[GT / 32];
  const int slot = blockIdx.x, tid = threadIdx.x;
  const int n = P.outN[slot];
  const size_t base = (size_t)slot * P.kpStride;
  const orbx_keypoint* in = P.outKp + base;
  orbx_keypoint* un = P.keysUn + base;
  u32* gOff = P.gridOff + (size_t)slot * (kGridCells + 1);
  u32* gIdx = P.gridIdx + base;
  for (int c = tid; c < kGridCells; c += GT) sCnt[c] = 0;
  __syncthreads();
  const float minX = (float)P.bounds[0], minY = (float)P.bounds[2];
  for (int i = tid; i < n; i += GT) {
    orbx_keypoint k = in[i];
    if (P.undistort) undistort_point(P, k.x, k.y, &k.x, &k.y);
    un[i] = k;
    const int gx = (int)roundf(__fmul_rn(k.x - minX, P.gridWInv));
    const int gy = (int)roundf(__fmul_rn(k.y - minY, P.gridHInv));
    if (gx >= 0 && gx < kGridCols && gy >= 0 && gy < kGridRows) atomicAdd(&sCnt[gx * kGridRows + gy], 1);
  }
  __syncthreads();
  // exclusive scan of the cell counts
  int loc[kCellsPerThread], s = 0;
#pragma unroll
  for (int j = 0; j < kCellsPerThread; ++j) { loc[j] = sCnt[tid * kCellsPerThread + j]; s += loc[j]; }
  int all;
  int run = cta_prefix_sum<GT>(s, sPart, &all) - s;
#pragma unroll
  for (int j = 0; j < kCellsPerThread; ++j) {
    const int c = tid * kCellsPerThread + j;
    gOff[c] = (u32)run;
    sCnt[c] = run;  // becomes the fill cursor
    run += loc[j];
  }
  if (tid == GT - 1) gOff[kGridCells] = (u32)run;
  __syncthreads();
  for (int i = tid; i < n; i += GT) {
    const orbx_keypoint k = un[i];
    const int gx = (int)roundf(__fmul_rn(k.x - minX, P.gridWInv));
    const int gy = (int)roundf(__fmul_rn(k.y - minY, P.gridHInv));
    if (gx >= 0 && gx < kGridCols && gy >= 0 && gy < kGridRows)
      gIdx[atomicAdd(&sCnt[gx * kGridRows + gy], 1)] = (u32)i;
  }
  __syncthreads();
  // ascending order inside each cell (lists are a handful of entries)
  for (int c = tid; c < kGridCells; c += GT) {
    const int b = (int)gOff[c], e = sCnt[c];
    for (int i = b + 1; i < e; ++i) {
      const u32 v = gIdx[i];
      int j = i - 1;
      while (j >= b && gIdx[j] > v) { gIdx[j + 1] = gIdx[j]; --j; }
      gIdx[j + 1] = v;
    }
  }
}

int launch_glue(const Params& P, int nSlots, cudaStream_t st) {
  if (nSlots <= 8) k_glue<1024><<<nSlots, 1024, 0, st>>>(P);
  else k_glue<256><<<nSlots, 256, 0, st>>>(P);
  return 1;
}

}  // namespace orbx
