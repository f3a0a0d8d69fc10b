// Scratch probe (not product): variants of a TMA tile load, one per process (argv[1] = variant).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
__device__ __forceinline__ unsigned sa(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// variant 0: documented libcu++ path, 2D
__global__ void k_doc2d(const __grid_constant__ CUtensorMap m, int x, int y, unsigned char* out) {
  __shared__ alignas(128) unsigned char buf[44][64];
#pragma nv_diag_suppress static_var_with_dynamic_init
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token token;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_2d_global_to_shared(&buf, &m, x, y, bar);
    token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(buf));
  } else token = bar.arrive();
  bar.wait(std::move(token));
  for (int i = threadIdx.x; i < 44 * 64; i += blockDim.x) out[i] = (&buf[0][0])[i];
}
// variant 1/2: raw PTX, 2D / 3D
template <int RANK>
__global__ void k_raw(const __grid_constant__ CUtensorMap m, int x, int y, int z, unsigned char* out) {
  __shared__ alignas(128) unsigned char buf[44 * 64];
  __shared__ alignas(8) unsigned long long bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(&bar)));
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(&bar)), "r"(44 * 64) : "memory");
    if (RANK == 2)
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                   ::"r"(sa(buf)), "l"((unsigned long long)&m), "r"(x), "r"(y), "r"(sa(&bar)) : "memory");
    else
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                   ::"r"(sa(buf)), "l"((unsigned long long)&m), "r"(x), "r"(y), "r"(z), "r"(sa(&bar)) : "memory");
  }
  unsigned done = 0;
  while (!done)
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(sa(&bar)) : "memory");
  for (int i = threadIdx.x; i < 44 * 64; i += blockDim.x) out[i] = buf[i];
}
int main(int argc, char** argv) {
  const int variant = argc > 1 ? atoi(argv[1]) : 0;
  const int W = 640, H = 480, pitch = 640, B = 4, bw = 64, bh = 44;
  size_t slab = (size_t)pitch * 520;
  unsigned char* d; cudaMalloc(&d, slab * B + 256);
  std::vector<unsigned char> h(slab * B);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)((i * 2654435761u) >> 13);
  cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  const int rank = (variant == 2 || variant == 3) ? 3 : 2;
  alignas(64) CUtensorMap m;
  cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)slab};
  cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
  CUresult r = ((EncodeFn)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, variant == 3 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("variant %d rank %d encode %d: ", variant, rank, (int)r);
  unsigned char* o; cudaMalloc(&o, bw * bh);
  const int x = argc > 2 ? atoi(argv[2]) : 16, y = 17, z = rank == 3 ? 2 : 0;
  if (variant == 0) k_doc2d<<<1, 128>>>(m, x, y, o);
  else if (variant == 1) k_raw<2><<<1, 128>>>(m, x, y, 0, o);
  else k_raw<3><<<1, 128>>>(m, x, y, z, o);
  cudaError_t er = cudaDeviceSynchronize();
  printf("%s", cudaGetErrorString(er));
  if (er == cudaSuccess) {
    std::vector<unsigned char> ho(bw * bh);
    cudaMemcpy(ho.data(), o, bw * bh, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < bh; ++rr) for (int c = 0; c < bw; ++c) bad += h[(size_t)z * slab + (size_t)(y + rr) * pitch + x + c] != ho[rr * bw + c];
    printf(" mismatches %d", bad);
  }
  printf("\n");
  return 0;
}
