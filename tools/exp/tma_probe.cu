// Scratch probe (not product): which way of handing a tensor map to cp.async.bulk.tensor works on this box.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <vector>
struct Maps { CUtensorMap m[16]; };
struct Big { int a[462]; };
__device__ __forceinline__ unsigned sa(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ void tma3(unsigned char* dst, unsigned long long* bar, const CUtensorMap* tm, int x, int y, int z, int bytes) {
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(sa(dst)), "l"((unsigned long long)tm), "r"(x), "r"(y), "r"(z), "r"(sa(bar)) : "memory");
  }
  __syncthreads();
  unsigned done = 0;
  while (!done)
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(sa(bar)) : "memory");
}
__global__ void kA(const __grid_constant__ CUtensorMap m, int x, int y, int z, int bw, int bh, unsigned char* out) {
  extern __shared__ __align__(128) unsigned char buf[];
  __shared__ __align__(8) unsigned long long bar;
  tma3(buf, &bar, &m, x, y, z, bw * bh);
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = buf[i];
}
__global__ void kB(const __grid_constant__ Big big, const __grid_constant__ Maps M, int lvl, int x, int y, int z, int bw, int bh, unsigned char* out) {
  extern __shared__ __align__(128) unsigned char buf[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ int extra[6];
  if (threadIdx.x == 0) extra[0] = big.a[lvl];
  tma3(buf, &bar, &M.m[lvl], x, y, z, bw * bh);
  for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = buf[i] + (extra[0] & 0);
}
int main() {
  const int W = 640, H = 480, pitch = 640, B = 4, bw = 64, bh = 44;
  size_t slab = (size_t)pitch * H + 256 * 100;
  unsigned char* d; cudaMalloc(&d, slab * B + 256);
  std::vector<unsigned char> h(slab * B);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)((i * 2654435761u) >> 13);
  cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  printf("entry: %d %d %p\n", (int)e, (int)q, fn);
  Maps M; memset(&M, 0, sizeof(M));
  for (int l = 0; l < 2; ++l) {
    cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)slab};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
    CUresult r = ((EncodeFn)fn)(&M.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d: %d\n", l, (int)r);
  }
  unsigned char* o; cudaMalloc(&o, bw * bh);
  std::vector<unsigned char> ho(bw * bh);
  auto check = [&](const char* name, int x, int y, int z) {
    cudaError_t er = cudaDeviceSynchronize();
    printf("%s: %s", name, cudaGetErrorString(er));
    if (er == cudaSuccess) {
      cudaMemcpy(ho.data(), o, bw * bh, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int r = 0; r < bh; ++r) for (int c = 0; c < bw; ++c) {
        int gx = x + c, gy = y + r;
        unsigned char ex = (gx < 0 || gx >= W || gy < 0 || gy >= H) ? 0 : h[(size_t)z * slab + (size_t)gy * pitch + gx];
        bad += ex != ho[r * bw + c];
      }
      printf(" mismatches %d", bad);
    }
    printf("\n");
  };
  kA<<<1, 128, bw * bh + 64>>>(M.m[0], 13, 17, 2, bw, bh, o); check("A direct param", 13, 17, 2);
  Big big; memset(&big, 0, sizeof(big));
  kB<<<1, 128, bw * bh + 64>>>(big, M, 1, 600, 450, 3, bw, bh, o); check("B struct array param, OOB corner", 600, 450, 3);
  return 0;
}
