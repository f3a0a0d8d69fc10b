// orbx_hamming_best2_noalloc: the raw all-pairs Hamming best / second best (orbx_hamming_best2) for callers that cannot call
// cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid allocation.  No counterpart in the
// reference.
//
// The existing entry point allocates its staging buffers and the per-query table k_hamming_topk writes (8 words per query:
// the 7 smallest (distance << 20 | index) keys, sorted, and an info word).  Here the caller passes DEVICE inputs and outputs,
// and the table lives in a file-scope device array of fixed size: kNoAllocMaxQueries rows = 512 KiB, held by every GPU that
// loads the library until the process exits.  All calls share that array: two calls that overlap on the GPU (different
// streams, or one GPU driven from several host threads) race on it.  Nothing in it has to start at zero -- k_hamming_topk
// writes all 8 words of every row below n_a before k_rows_to_best2 reads them -- so it is never cleared.
//
// In host code `g_rows` is only a stand-in for the device array: its device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and that pointer is what the kernels get.
#include <mutex>

#include "ctx.cuh"

namespace orbx {

constexpr int kNoAllocMaxQueries = 16384;  // rows of the table (ORBX_NOALLOC_MAX_QUERIES in orbx.h)
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) u32 g_rows[(size_t)kNoAllocMaxQueries * 8];

static std::mutex g_rowsMutex;
static u32* g_rowsPtr[kNoAllocMaxDevices];  // device address of g_rows per device ordinal (nullptr: not looked up yet)

// best index / best distance / second-best distance of each query from its sorted row (as the host loop of orbx_hamming_best2)
__global__ void __launch_bounds__(256) k_rows_to_best2(const u32* __restrict__ rows, int nA, int* bestIdx, int* bestDist, int* secondDist) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nA) return;
  const uint2 k = __ldg(reinterpret_cast<const uint2*>(rows + (size_t)q * 8));
  if (bestIdx) bestIdx[q] = (int)(k.x & 0xfffffu);
  if (bestDist) bestDist[q] = (int)(k.x >> 20);
  if (secondDist) secondDist[q] = k.y == 0xffffffffu ? -1 : (int)(k.y >> 20);
}

// the device address of g_rows on the CURRENT device; looked up once per device
static cudaError_t rows_ptr(int device, u32** out) {
  std::lock_guard<std::mutex> lock(g_rowsMutex);
  if (!g_rowsPtr[device]) {
    u32* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_rows);
    if (e != cudaSuccess) return e;
    g_rowsPtr[device] = p;
  }
  *out = g_rowsPtr[device];
  return cudaSuccess;
}

}  // namespace orbx

using namespace orbx;

extern "C" int orbx_hamming_best2_noalloc(orbx_ctx* c, const uint8_t* d_a, int n_a, const uint8_t* d_b, int n_b, int* d_best_idx,
                                          int* d_best_dist, int* d_second_dist, void* stream, int synchronize) {
  if (!c) return ORBX_ERR_ARG;
  if (n_a < 1 || n_b < 1 || n_b >= (1 << 20) || !d_a || !d_b) {
    c->err = "orbx_hamming_best2_noalloc: null input or n_a < 1 or n_b outside [1, 2^20)";
    return ORBX_ERR_ARG;
  }
  if (n_a > kNoAllocMaxQueries) {
    c->err = "orbx_hamming_best2_noalloc: n_a = " + std::to_string(n_a) + " exceeds the fixed scratch table of " +
             std::to_string(kNoAllocMaxQueries) + " query rows (ORBX_NOALLOC_MAX_QUERIES); split the queries or use orbx_hamming_best2";
    return ORBX_ERR_CAPACITY;
  }
  if (c->device < 0 || c->device >= kNoAllocMaxDevices) { c->err = "orbx_hamming_best2_noalloc: device ordinal out of range"; return ORBX_ERR_ARG; }
  CU(c, cudaSetDevice(c->device));
  u32* rows = nullptr;
  CU(c, rows_ptr(c->device, &rows));
  cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
  launch_hamming_topk(d_a, n_a, d_b, n_b, rows, st);
  k_rows_to_best2<<<(n_a + 255) / 256, 256, 0, st>>>(rows, n_a, d_best_idx, d_best_dist, d_second_dist);
  CU(c, cudaGetLastError());
  c->launches = 2;
  if (synchronize) CU(c, cudaStreamSynchronize(st));
  return ORBX_OK;
}
