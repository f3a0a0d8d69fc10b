// Probe: small-box TMA copies (W words x R rows x 1) per warp into shared memory, to find which issue pattern is legal/fast.
// usage: tma_patch_probe MODE   (one mode per process: a fault is sticky)
//   0: thread 0 of the CTA issues the copies of all 8 warps      1: lane 0 of every warp issues its own
//   2: like 1, tensor map read from global memory                 3: like 1, box 12 words wide
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
typedef unsigned int u32; typedef unsigned long long u64;
__device__ __forceinline__ u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }
struct Maps { CUtensorMap m[2]; };
constexpr int PER = 4;
static int g_xmask = ~0;
template <int W, int R, int MODE, bool DYN = false>
__global__ void __launch_bounds__(256) k(const __grid_constant__ Maps P, const Maps* G, const int* xs, const int* ys, u32* out, int n) {
  constexpr int BYTES = W * R * 4, BUF = (BYTES + 127) & ~127;
  __shared__ __align__(128) u32 sPs[DYN ? 1 : 8][2][BUF / 4];
  extern __shared__ __align__(128) u32 dyn[];
  u32 (*sP)[2][BUF / 4] = DYN ? reinterpret_cast<u32 (*)[2][BUF / 4]>(dyn) : sPs;
  __shared__ __align__(8) u64 sBar[8][2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int w = 0; w < 8; ++w) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sBar[w][0])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sBar[w][1])));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  auto issue_for = [&](int w, int i) {
    const int base = (blockIdx.x * 8 + w) * PER;
    if (base + i >= n) return;
    const u32 bar = smem_addr(&sBar[w][i & 1]);
    const CUtensorMap* map = MODE == 2 ? &G->m[i & 1] : &P.m[i & 1];
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(BYTES) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(smem_addr(&sP[w][i & 1][0])), "l"(reinterpret_cast<u64>(map)), "r"(xs[base + i] >> 2), "r"(ys[base + i]), "r"(0), "r"(bar) : "memory");
  };
  auto issue = [&](int i) {
    if (MODE == 0) { if (tid == 0) for (int w = 0; w < 8; ++w) issue_for(w, i); }
    else if (lane == 0) issue_for(warp, i);
  };
  issue(0); issue(1);
  const int base = (blockIdx.x * 8 + warp) * PER;
  u32 acc = 0;
  for (int i = 0; i < PER; ++i) {
    if (base + i < n) {
      const u32 bar = smem_addr(&sBar[warp][i & 1]), parity = (u32)(i >> 1) & 1u;
      u32 done = 0;
      while (!done) asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
      for (int j = lane; j < W * R; j += 32) acc += sP[warp][i & 1][j];
    }
    if (MODE == 0) __syncthreads(); else __syncwarp();
    if (i + 2 < PER) issue(i + 2);
  }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[blockIdx.x * 8 + warp] = acc;
}
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
template <int W, int R, int MODE, bool DYN = false> int run(void* fn, unsigned char* img, const std::vector<unsigned char>& himg, int pitch, int h, int n, const char* name) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                               CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  Maps M;
  cuuint64_t dims[3] = {(cuuint64_t)pitch / 4, (cuuint64_t)h, 1}; cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * h};
  cuuint32_t box[3] = {W, R, 1}, es[3] = {1, 1, 1};
  for (int b = 0; b < 2; ++b) {
    CUresult r = ((EncodeFn)fn)(&M.m[b], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, img, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("%s: encode failed %d\n", name, (int)r); return 1; }
  }
  Maps* dM; CK(cudaMalloc(&dM, sizeof(Maps))); CK(cudaMemcpy(dM, &M, sizeof(Maps), cudaMemcpyHostToDevice));
  std::vector<int> xs(n), ys(n);
  for (int i = 0; i < n; ++i) { xs[i] = (1 + rand() % (pitch - 70)) & g_xmask; ys[i] = 1 + rand() % (h - 40); }
  int *dx, *dy; u32* dout; CK(cudaMalloc(&dx, n * 4)); CK(cudaMalloc(&dy, n * 4)); CK(cudaMalloc(&dout, (n / PER + 8) * 4));
  CK(cudaMemcpy(dx, xs.data(), n * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dy, ys.data(), n * 4, cudaMemcpyHostToDevice));
  const int grid = (n + 8 * PER - 1) / (8 * PER);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<W, R, MODE, DYN><<<grid, 256, DYN ? 16 * ((W * R * 4 + 127) & ~127) : 0>>>(M, dM, dx, dy, dout, n);
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  for (int it = 0; it < 5; ++it) k<W, R, MODE, DYN><<<grid, 256, DYN ? 16 * ((W * R * 4 + 127) & ~127) : 0>>>(M, dM, dx, dy, dout, n);
  cudaEventRecord(e1); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<u32> out(n / PER); CK(cudaMemcpy(out.data(), dout, (n / PER) * 4, cudaMemcpyDeviceToHost));
  u32 ref = 0;
  for (int i = 0; i < PER; ++i) for (int r = 0; r < R; ++r) for (int w = 0; w < W; ++w) {
    const size_t o = (size_t)(ys[i] + r) * pitch + (size_t)((xs[i] >> 2) + w) * 4;
    ref += *(const u32*)&himg[o];
  }
  printf("%s: %.3f ms per launch of %d boxes (%d x %d words) = %.2f G boxes/s, %.1f GB/s; check %s\n", name, ms / 5, n, W, R, n / (ms / 5) / 1e6,
         (double)n * W * R * 4 / (ms / 5) / 1e6, ref == out[0] ? "ok" : "MISMATCH");
  return 0;
}
int main(int argc, char** argv) {
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  if (argc > 2) g_xmask = ~(atoi(argv[2]) - 1);  // x coordinates (bytes) aligned to this
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaFree(0));
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  const int pitch = 640, h = 480 * 256, n = 1 << 20;
  unsigned char* img; CK(cudaMalloc(&img, (size_t)pitch * h));
  std::vector<unsigned char> hi((size_t)pitch * h); for (auto& b : hi) b = rand();
  CK(cudaMemcpy(img, hi.data(), hi.size(), cudaMemcpyHostToDevice));
  if (mode == 0) return run<16, 37, 0>(fn, img, hi, pitch, h, n, "cta thread 0, 16x37");
  if (mode == 1) return run<16, 37, 1>(fn, img, hi, pitch, h, n, "lane 0 per warp, 16x37");
  if (mode == 2) return run<16, 37, 2>(fn, img, hi, pitch, h, n, "lane 0 per warp, map in global, 16x37");
  if (mode == 3) return run<12, 37, 1>(fn, img, hi, pitch, h, n, "lane 0 per warp, 12x37");
  if (mode == 4) return run<16, 37, 1, true>(fn, img, hi, pitch, h, n, "lane 0 per warp, dynamic smem, 16x37");
  if (mode == 5) return run<16, 32, 1>(fn, img, hi, pitch, h, n, "lane 0 per warp, 16x32");
  if (mode == 6) return run<64, 8, 1>(fn, img, hi, pitch, h, n, "lane 0 per warp, 64x8");
  if (mode == 7) return run<64, 8, 0, true>(fn, img, hi, pitch, h, n, "cta thread 0, dynamic, 64x8");
  return 2;
}
