// C ABI, part 3: camera / Frame glue / matcher entry points on host frames, and the
// device-resident stream path (extract -> glue -> SearchForInitialization without leaving HBM).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "ctx.cuh"

using namespace orbx;

namespace orbx {
void launch_features_in_area(const Params& P, const orbx_keypoint* kun, const u32* gOff, const u32* gIdx, float x,
                             float y, float r, int minLevel, int maxLevel, int* out, int cap, int* count,
                             cudaStream_t st);
void launch_features_in_area_batch(const Params& P, const orbx_keypoint* keysUn, const u32* gridOff, const u32* gridIdx, int stride,
                                   const int* slots, const float* q5, int nQ, int* out, int cap, int* counts, cudaStream_t st);
}

static int score_reserve(orbx_ctx* c, size_t bytes);  // scratch block shared by the scoring and the batched query calls

// cv::undistortPoints for the four image corners (ComputeImageBounds, SlamTypes/Frame.cpp:101-134).
// Runs once per context on the host: it is constructor arithmetic like the level tables, and its
// float -> int truncation decides the grid geometry, so it is evaluated exactly as the CPU does.
static void undistort_host(float u, float v, const float K[4], const float D[4], float* ou, float* ov) {
  const double fx = K[0], fy = K[1], cx = K[2], cy = K[3], k1 = D[0], k2 = D[1], p1 = D[2], p2 = D[3];
  const double ifx = 1. / fx, ify = 1. / fy;
  double x = ((double)u - cx) * ifx, y = ((double)v - cy) * ify;
  const double x0 = x, y0 = y;
  for (int j = 0; j < 5; ++j) {
    const double r2 = x * x + y * y;
    const double icdist = 1. / (1 + ((0 * r2 + k2) * r2 + k1) * r2);
    if (icdist < 0) { x = x0; y = y0; break; }  // OpenCV 4.x: negative radial factor -> the start value
    const double dx = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
    const double dy = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
    x = (x0 - dx) * icdist;
    y = (y0 - dy) * icdist;
  }
  *ou = (float)(x * fx + cx);
  *ov = (float)(y * fy + cy);
}

extern "C" int orbx_set_camera(orbx_ctx* c, const float K4[4], const float dist4[4], const int bounds[4],
                               int bounds_out[4]) {
  if (!c || !K4 || !dist4) return ORBX_ERR_ARG;
  Params& P = c->P;
  std::memcpy(P.K4, K4, sizeof(P.K4));
  std::memcpy(P.D4, dist4, sizeof(P.D4));
  P.undistort = dist4[0] != 0.0f;  // :138, :103
  if (bounds) {
    std::memcpy(P.bounds, bounds, sizeof(P.bounds));
  } else if (P.undistort) {
    const float cs[4][2] = {{0.f, 0.f}, {(float)P.W, 0.f}, {0.f, (float)P.H}, {(float)P.W, (float)P.H}};
    float u[4][2];
    for (int i = 0; i < 4; ++i) undistort_host(cs[i][0], cs[i][1], K4, dist4, &u[i][0], &u[i][1]);
    P.bounds[0] = (int)std::min(u[0][0], u[2][0]);
    P.bounds[1] = (int)std::max(u[1][0], u[3][0]);
    P.bounds[2] = (int)std::min(u[0][1], u[1][1]);
    P.bounds[3] = (int)std::max(u[2][1], u[3][1]);
  } else {
    P.bounds[0] = 0; P.bounds[1] = P.W; P.bounds[2] = 0; P.bounds[3] = P.H;
  }
  if (P.bounds[1] <= P.bounds[0] || P.bounds[3] <= P.bounds[2]) return ORBX_ERR_ARG;
  P.gridWInv = (float)kGridCols / (float)(P.bounds[1] - P.bounds[0]);  // :46-47
  P.gridHInv = (float)kGridRows / (float)(P.bounds[3] - P.bounds[2]);
  if (bounds_out) std::memcpy(bounds_out, P.bounds, sizeof(P.bounds));
  c->cameraSet = 1;
  ++c->graphEpoch;
  return ORBX_OK;
}

extern "C" int orbx_frame_glue(orbx_ctx* c, const orbx_keypoint* kps, int n, orbx_keypoint* kps_un,
                               uint32_t* grid_offsets, uint32_t* grid_indices) {
  if (!c || n < 0 || (n > 0 && !kps)) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  const Params& P = c->P;
  if (n > P.kpStride) return ORBX_ERR_CAPACITY;
  CU(c, cudaSetDevice(c->device));
  { const int rc = drain_flights(c); if (rc) return rc; }
  cudaStream_t st = c->stream;
  // slot 0 of the frame store is the staging area of this call
  if (n) CU(c, cudaMemcpyAsync(P.outKp, kps, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, st));
  CU(c, cudaMemcpyAsync(P.outN, &n, sizeof(int), cudaMemcpyHostToDevice, st));
  c->launches = launch_glue(P, 1, st);
  CU(c, cudaGetLastError());
  if (kps_un && n) CU(c, cudaMemcpyAsync(kps_un, P.keysUn, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st));
  if (grid_offsets)
    CU(c, cudaMemcpyAsync(grid_offsets, P.gridOff, (kGridCells + 1) * sizeof(u32), cudaMemcpyDeviceToHost, st));
  if (grid_indices && n) CU(c, cudaMemcpyAsync(grid_indices, P.gridIdx, (size_t)n * sizeof(u32), cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  return ORBX_OK;
}

// copies a host frame into staging slot s (0 or 1) of the PINNED mirror of the staging store; stage_commit() sends the mirror to
// the device in one copy (the caller's vectors are pageable: one asynchronous copy per array cost ~10 us each, ten per pair)
static int stage_frame(orbx_ctx* c, const orbx_frame* F, int s) {
  if (!F || F->n < 0) return ORBX_ERR_ARG;
  if (F->n > c->hfCap) return ORBX_ERR_CAPACITY;
  if (F->n > 0 && (!F->keys_un || !F->desc || !F->grid_indices)) return ORBX_ERR_ARG;
  if (!F->grid_offsets) return ORBX_ERR_ARG;
  const size_t cap = (size_t)c->hfCap;
  u8* const h = c->hHfIn;
  auto at = [&](const void* dev) { return h + (reinterpret_cast<const u8*>(dev) - c->hfBlock); };
  if (F->n) {
    const u32 nIdx = F->grid_offsets[kGridCells];
    if (nIdx > (u32)F->n) return ORBX_ERR_ARG;
    std::memcpy(at(c->hfKeys + s * cap), F->keys_un, (size_t)F->n * sizeof(orbx_keypoint));
    std::memcpy(at(c->hfDesc + s * cap * 32), F->desc, (size_t)F->n * 32);
    if (nIdx) std::memcpy(at(c->hfGridIdx + s * cap), F->grid_indices, nIdx * sizeof(u32));
  }
  std::memcpy(at(c->hfGridOff + (size_t)s * (kGridCells + 1)), F->grid_offsets, (kGridCells + 1) * sizeof(u32));
  std::memcpy(at(c->hfN + s), &F->n, sizeof(int));
  return ORBX_OK;
}
static int stage_commit(orbx_ctx* c) {
  CU(c, cudaMemcpyAsync(c->hfBlock, c->hHfIn, c->hfInBytes, cudaMemcpyHostToDevice, c->stream));
  return ORBX_OK;
}
// the results of a host-frame matcher call (nmatches | counters | matches12[n1]) through the pinned mirror: one D2H copy, one sync
static int fetch_match_results(orbx_ctx* c, int n1, int* matches12, int* nmatches, int* counters) {
  const size_t bytes = 16 + (matches12 ? (size_t)n1 * sizeof(int) : 0);
  CU(c, cudaMemcpyAsync(c->hHfOut, c->hfBlock + c->hfInBytes, bytes, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  if (nmatches) std::memcpy(nmatches, c->hHfOut, sizeof(int));
  if (counters) std::memcpy(counters, c->hHfOut + 4, 3 * sizeof(int));
  if (matches12 && n1) std::memcpy(matches12, c->hHfOut + 16, (size_t)n1 * sizeof(int));
  return ORBX_OK;
}

extern "C" int orbx_features_in_area(orbx_ctx* c, const orbx_frame* F, float x, float y, float r, int min_level,
                                     int max_level, int* out, int cap) {
  if (!c || !F || cap < 0 || (cap > 0 && !out)) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  CU(c, cudaSetDevice(c->device));
  int rc = drain_flights(c);
  if (rc) return rc;
  if ((rc = stage_frame(c, F, 0)) || (rc = stage_commit(c))) return rc;
  const int dcap = std::min(cap, c->hfCap);
  launch_features_in_area(c->P, c->hfKeys, c->hfGridOff, c->hfGridIdx, x, y, r, min_level, max_level, c->hfMatches,
                          dcap, c->hfNMatches, c->stream);
  CU(c, cudaGetLastError());
  int n = 0;
  CU(c, cudaMemcpyAsync(&n, c->hfNMatches, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  const int m = std::min(n, dcap);
  if (m > 0) CU(c, cudaMemcpy(out, c->hfMatches, (size_t)m * sizeof(int), cudaMemcpyDeviceToHost));
  return n;
}

// Radius queries on the frames the stream path left on the device: no frame is uploaded; n queries in one
// launch, a warp each.
extern "C" int orbx_stream_features_in_area(orbx_ctx* c, int n_queries, const int* slots, const float* queries5, int* out, int cap,
                                            int* counts) {
  if (!c || n_queries < 0 || cap < 0) return ORBX_ERR_ARG;
  if (n_queries == 0) return ORBX_OK;
  if (!slots || !queries5 || !counts || (cap > 0 && !out)) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  for (int i = 0; i < n_queries; ++i)
    if (slots[i] < 0 || slots[i] >= c->P.maxBatch) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  int rc = drain_flights(c);
  if (rc) return rc;
  const size_t nq = (size_t)n_queries, need = nq * sizeof(int) * 2 + nq * 5 * sizeof(float) + nq * (size_t)std::max(cap, 1) * sizeof(int) + 1024;
  if ((rc = score_reserve(c, need))) return rc;   // the scratch block of the "next row" calls
  u8* base = c->dScore;
  int* dSlots = reinterpret_cast<int*>(base);
  int* dCounts = dSlots + nq;
  float* dQ = reinterpret_cast<float*>(dCounts + nq);
  int* dOut = reinterpret_cast<int*>(dQ + 5 * nq);
  cudaStream_t st = c->stream;
  CU(c, cudaMemcpyAsync(dSlots, slots, nq * sizeof(int), cudaMemcpyHostToDevice, st));
  CU(c, cudaMemcpyAsync(dQ, queries5, nq * 5 * sizeof(float), cudaMemcpyHostToDevice, st));
  launch_features_in_area_batch(c->P, c->P.keysUn, c->P.gridOff, c->P.gridIdx, c->P.kpStride, dSlots, dQ, n_queries, dOut, cap, dCounts, st);
  c->launches = 1;
  CU(c, cudaGetLastError());
  CU(c, cudaMemcpyAsync(counts, dCounts, nq * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (cap > 0) CU(c, cudaMemcpyAsync(out, dOut, nq * (size_t)cap * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  return ORBX_OK;
}

extern "C" int orbx_search_for_initialization(orbx_ctx* c, const orbx_frame* F1, const orbx_frame* F2, float nnratio,
                                              int check_orientation, int window_size, int* matches12, int* nmatches,
                                              int counters[3]) {
  if (!c || !F1 || !F2) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  CU(c, cudaSetDevice(c->device));
  int rc = drain_flights(c);
  if (rc) return rc;
  if ((rc = stage_frame(c, F1, 0))) return rc;
  if ((rc = stage_frame(c, F2, 1)) || (rc = stage_commit(c))) return rc;
  MatchArgs A;
  A.keysUn = c->hfKeys; A.desc = c->hfDesc; A.gridOff = c->hfGridOff; A.gridIdx = c->hfGridIdx; A.n = c->hfN;
  A.stride = c->hfCap; A.matches12 = c->hfMatches; A.nmatches = c->hfNMatches; A.counters = c->hfCounters;
  A.scratch = c->hfScratch;
  A.bf = c->hfBf;
  c->launches = launch_match(c->P, A, 1, 2, nnratio, check_orientation, window_size, c->stream);
  CU(c, cudaGetLastError());
  return fetch_match_results(c, F1->n, matches12, nmatches, counters);
}

// ---- the distance matrix of one pair sharded by query rows (multi-GPU brute-force matching) ------------------------
static MatchArgs host_pair_args(orbx_ctx* c) {
  MatchArgs A;
  A.keysUn = c->hfKeys; A.desc = c->hfDesc; A.gridOff = c->hfGridOff; A.gridIdx = c->hfGridIdx; A.n = c->hfN;
  A.stride = c->hfCap; A.matches12 = c->hfMatches; A.nmatches = c->hfNMatches; A.counters = c->hfCounters;
  A.scratch = c->hfScratch;
  A.bf = c->hfBf;
  return A;
}

extern "C" int orbx_match_rows(orbx_ctx* c, const orbx_frame* F1, const orbx_frame* F2, int window_size, int q0, int q1,
                               uint32_t* rows_out, int* n_queries) {
  if (!c || !F1 || !F2 || q0 < 0 || q1 < q0) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  if (c->hfCap >= 45000) return ORBX_ERR_CAPACITY;
  CU(c, cudaSetDevice(c->device));
  int rc = stage_frame(c, F1, 0);
  if (rc) return rc;
  if ((rc = stage_frame(c, F2, 1)) || (rc = stage_commit(c))) return rc;
  const MatchArgs A = host_pair_args(c);
  c->launches = launch_bf_prep(c->P, A, A.bf, 1, 2, c->stream);
  int cnt[4] = {0, 0, 0, 0};
  CU(c, cudaMemcpyAsync(cnt, A.bf.counts, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  if (n_queries) *n_queries = cnt[1];
  q1 = std::min(q1, cnt[1]);
  if (q1 > q0) {
    c->launches += launch_bf_rows(c->P, A, A.bf, 1, 2, window_size, q0, q1, c->stream);
    CU(c, cudaGetLastError());
    if (rows_out)
      CU(c, cudaMemcpyAsync(rows_out, A.bf.topk + (size_t)q0 * 8, (size_t)(q1 - q0) * 8 * sizeof(u32), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
  }
  return std::max(0, q1 - q0);
}

extern "C" int orbx_match_resolve(orbx_ctx* c, const orbx_frame* F1, const orbx_frame* F2, float nnratio, int check_orientation,
                                  int window_size, const uint32_t* rows, int n_queries, int* matches12, int* nmatches,
                                  int counters[3]) {
  if (!c || !F1 || !F2 || n_queries < 0 || (n_queries > 0 && !rows)) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  if (c->hfCap >= 45000 || n_queries > c->hfCap) return ORBX_ERR_CAPACITY;
  CU(c, cudaSetDevice(c->device));
  int rc = stage_frame(c, F1, 0);
  if (rc) return rc;
  if ((rc = stage_frame(c, F2, 1)) || (rc = stage_commit(c))) return rc;
  const MatchArgs A = host_pair_args(c);
  c->launches = launch_bf_prep(c->P, A, A.bf, 1, 2, c->stream);
  if (n_queries) CU(c, cudaMemcpyAsync(A.bf.topk, rows, (size_t)n_queries * 8 * sizeof(u32), cudaMemcpyHostToDevice, c->stream));
  int cnt[4] = {0, 0, 0, 0};
  CU(c, cudaMemcpyAsync(cnt, A.bf.counts, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  if (cnt[1] != n_queries) return ORBX_ERR_ARG;  // the row table does not belong to this pair
  c->launches += launch_bf_resolve(c->P, A, A.bf, 1, 2, nnratio, check_orientation, window_size, c->stream);
  CU(c, cudaGetLastError());
  return fetch_match_results(c, F1->n, matches12, nmatches, counters);
}

// Raw all-pairs Hamming best / second best (micro-benchmark of the brute-force configuration; DescriptorDistance over a matrix)
extern "C" int orbx_hamming_best2(orbx_ctx* c, const uint8_t* a, int n_a, const uint8_t* b, int n_b, int* best_idx, int* best_dist,
                                  int* second_dist, int repeats, float* kernel_ms) {
  if (!c || n_a < 1 || n_b < 1 || n_b >= (1 << 20) || !a || !b) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  u8 *da = nullptr, *db = nullptr;
  u32* dt = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  int rc = ORBX_OK;
  std::vector<u32> rows((size_t)n_a * 8);
  auto fail = [&](cudaError_t e) { if (e != cudaSuccess && rc == ORBX_OK) { c->err = cudaGetErrorString(e); rc = ORBX_ERR_CUDA; } };
  fail(cudaMalloc(&da, (size_t)n_a * 32));
  fail(cudaMalloc(&db, (size_t)n_b * 32));
  fail(cudaMalloc(&dt, (size_t)n_a * 8 * sizeof(u32)));
  fail(cudaEventCreate(&e0));
  fail(cudaEventCreate(&e1));
  if (rc == ORBX_OK) {
    fail(cudaMemcpyAsync(da, a, (size_t)n_a * 32, cudaMemcpyHostToDevice, c->stream));
    fail(cudaMemcpyAsync(db, b, (size_t)n_b * 32, cudaMemcpyHostToDevice, c->stream));
    launch_hamming_topk(da, n_a, db, n_b, dt, c->stream);  // warm-up
    fail(cudaEventRecord(e0, c->stream));
    const int reps = std::max(1, repeats);
    for (int i = 0; i < reps; ++i) launch_hamming_topk(da, n_a, db, n_b, dt, c->stream);
    fail(cudaEventRecord(e1, c->stream));
    fail(cudaMemcpyAsync(rows.data(), dt, rows.size() * sizeof(u32), cudaMemcpyDeviceToHost, c->stream));
    fail(cudaStreamSynchronize(c->stream));
    c->launches = reps + 1;
    float ms = 0.f;
    if (rc == ORBX_OK) fail(cudaEventElapsedTime(&ms, e0, e1));
    if (kernel_ms) *kernel_ms = ms / reps;
  }
  if (rc == ORBX_OK)
    for (int q = 0; q < n_a; ++q) {
      const u32 k0 = rows[(size_t)q * 8], k1 = rows[(size_t)q * 8 + 1];
      if (best_idx) best_idx[q] = (int)(k0 & 0xfffffu);
      if (best_dist) best_dist[q] = (int)(k0 >> 20);
      if (second_dist) second_dist[q] = k1 == 0xffffffffu ? -1 : (int)(k1 >> 20);
    }
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(da); cudaFree(db); cudaFree(dt);
  return rc;
}

extern "C" int orbx_descriptor_distance_batch(orbx_ctx* c, const uint8_t* a, const uint8_t* b, int n, int* out) {
  if (!c || n < 0 || (n > 0 && (!a || !b || !out))) return ORBX_ERR_ARG;
  if (n == 0) return ORBX_OK;
  CU(c, cudaSetDevice(c->device));
  u8 *da = nullptr, *db = nullptr;
  int* dout = nullptr;
  CU(c, cudaMalloc(&da, (size_t)n * 32));
  cudaError_t e1 = cudaMalloc(&db, (size_t)n * 32), e2 = cudaMalloc(&dout, (size_t)n * sizeof(int));
  int rc = ORBX_OK;
  if (e1 == cudaSuccess && e2 == cudaSuccess) {
    cudaMemcpyAsync(da, a, (size_t)n * 32, cudaMemcpyHostToDevice, c->stream);
    cudaMemcpyAsync(db, b, (size_t)n * 32, cudaMemcpyHostToDevice, c->stream);
    launch_hamming(da, db, n, dout, c->stream);
    cudaMemcpyAsync(out, dout, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, c->stream);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) { c->err = cudaGetErrorString(e); rc = ORBX_ERR_CUDA; }
  } else {
    c->err = "cudaMalloc failed"; rc = ORBX_ERR_CUDA;
  }
  cudaFree(da); cudaFree(db); cudaFree(dout);
  return rc;
}

// cos / sin of n float angles exactly as k_describe computes them in trig mode 2 (test hook of the restated glibc cosf / sinf)
extern "C" int orbx_sincosf_probe(orbx_ctx* c, const float* y, long long n, float* sn, float* cs) {
  if (!c || n < 0 || (n > 0 && (!y || !sn || !cs))) return ORBX_ERR_ARG;
  if (n == 0) return ORBX_OK;
  CU(c, cudaSetDevice(c->device));
  float* d = nullptr;
  CU(c, cudaMalloc(&d, (size_t)n * 3 * sizeof(float)));
  cudaMemcpyAsync(d, y, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, c->stream);
  launch_sincosf_probe(d, n, d + n, d + 2 * n, c->stream);
  cudaMemcpyAsync(sn, d + n, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
  cudaMemcpyAsync(cs, d + 2 * n, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
  const cudaError_t e = cudaStreamSynchronize(c->stream);
  cudaFree(d);
  if (e != cudaSuccess) { c->err = cudaGetErrorString(e); return ORBX_ERR_CUDA; }
  return ORBX_OK;
}

// ---- Initializer scoring -------------------------------------------------------------
static int score_reserve(orbx_ctx* c, size_t bytes) {
  if (bytes <= c->scoreBytes) return ORBX_OK;
  if (c->dScore) cudaFree(c->dScore);
  c->dScore = nullptr; c->scoreBytes = 0;
  CU(c, cudaMalloc(&c->dScore, bytes));
  c->scoreBytes = bytes;
  return ORBX_OK;
}
namespace {
size_t up256(size_t b) { return (b + 255) & ~(size_t)255; }
// invSigmaSquare = 1.0 / (sigma * sigma): float product, double division, narrowed (Initializer.cpp:293,376)
float inv_sigma_square(float sigma) { const float s2 = sigma * sigma; return (float)(1.0 / (double)s2); }

// one model kind over nPairs pairs whose frames / matches are already on the device; host in / out
int score_run(orbx_ctx* c, ScoreArgs S, int nPairs, bool fund, const float* hA, const float* hB, float* scores, int* best,
              int* nList, uint8_t* inlAll, int listLen, uint8_t* inlBest, int cap) {
  const size_t nh = (size_t)nPairs * S.nHyp;
  if ((size_t)S.stride * 16 + (size_t)S.nHyp * 4 > 200 * 1024) return ORBX_ERR_CAPACITY;
  const size_t oA = 0, oB = oA + up256(nh * 36), oS = oB + up256(nh * 36), oBest = oS + up256(nh * 4),
               oN = oBest + up256((size_t)nPairs * 4), oIB = oN + up256((size_t)nPairs * 4),
               oIA = oIB + up256((size_t)nPairs * S.stride), total = oIA + (inlAll ? up256(nh * (size_t)listLen) : 0);
  int rc = score_reserve(c, total);
  if (rc) return rc;
  cudaStream_t st = c->stream;
  u8* d = c->dScore;
  CU(c, cudaMemcpyAsync(d + oA, hA, nh * 36, cudaMemcpyHostToDevice, st));
  if (!fund) CU(c, cudaMemcpyAsync(d + oB, hB, nh * 36, cudaMemcpyHostToDevice, st));
  S.hypA = reinterpret_cast<const float*>(d + oA);
  S.hypB = reinterpret_cast<const float*>(d + oB);
  S.scores = reinterpret_cast<float*>(d + oS);
  S.best = reinterpret_cast<int*>(d + oBest);
  S.nList = reinterpret_cast<int*>(d + oN);
  S.inlBest = inlBest ? d + oIB : nullptr;
  S.inlAll = inlAll ? d + oIA : nullptr;
  if (c->stageTiming) cudaEventRecord(c->evStage[8], st);
  c->launches = launch_score_models(S, nPairs, fund, st);
  CU(c, cudaGetLastError());
  if (c->stageTiming) cudaEventRecord(c->evStage[9], st);
  if (scores) CU(c, cudaMemcpyAsync(scores, d + oS, nh * 4, cudaMemcpyDeviceToHost, st));
  if (best) CU(c, cudaMemcpyAsync(best, d + oBest, (size_t)nPairs * 4, cudaMemcpyDeviceToHost, st));
  if (nList) CU(c, cudaMemcpyAsync(nList, d + oN, (size_t)nPairs * 4, cudaMemcpyDeviceToHost, st));
  if (inlAll && nh * (size_t)listLen) CU(c, cudaMemcpyAsync(inlAll, d + oIA, nh * (size_t)listLen, cudaMemcpyDeviceToHost, st));
  if (inlBest)
    CU(c, cudaMemcpy2DAsync(inlBest, (size_t)cap, d + oIB, (size_t)S.stride, (size_t)std::min(cap, S.stride), nPairs,
                            cudaMemcpyDeviceToHost, st));
  CU(c, cudaStreamSynchronize(st));
  if (c->stageTiming) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->evStage[8], c->evStage[9]);
    c->stageMs[5] += ms;
  }
  return ORBX_OK;
}

// host frames of one pair -> staging slots 0 / 1; returns the list length in *listLen
int score_stage(orbx_ctx* c, const orbx_keypoint* k1, int n1, const orbx_keypoint* k2, int n2, const int* m12, int* listLen) {
  if (!c || n1 < 0 || n2 < 0 || (n1 > 0 && (!k1 || !m12)) || (n2 > 0 && !k2)) return ORBX_ERR_ARG;
  if (n1 > c->hfCap || n2 > c->hfCap) return ORBX_ERR_CAPACITY;
  int N = 0;
  for (int i = 0; i < n1; ++i) {
    if (m12[i] >= n2) return ORBX_ERR_ARG;
    N += m12[i] >= 0;
  }
  *listLen = N;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const size_t cap = (size_t)c->hfCap;
  const int n[2] = {n1, n2};
  if (n1) {
    CU(c, cudaMemcpyAsync(c->hfKeys, k1, (size_t)n1 * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, st));
    CU(c, cudaMemcpyAsync(c->hfMatches, m12, (size_t)n1 * sizeof(int), cudaMemcpyHostToDevice, st));
  }
  if (n2) CU(c, cudaMemcpyAsync(c->hfKeys + cap, k2, (size_t)n2 * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, st));
  CU(c, cudaMemcpyAsync(c->hfN, n, sizeof(n), cudaMemcpyHostToDevice, st));
  CU(c, cudaStreamSynchronize(st));  // n[] lives on this stack frame
  return ORBX_OK;
}
ScoreArgs host_pair_args(orbx_ctx* c, int nHyp, float sigma) {
  ScoreArgs S{};
  S.keysUn = c->hfKeys; S.n = c->hfN; S.stride = c->hfCap; S.pairStep = 2; S.matches12 = c->hfMatches;
  S.nHyp = nHyp; S.invSigmaSquare = inv_sigma_square(sigma);
  return S;
}
}  // namespace

extern "C" int orbx_check_homography(orbx_ctx* c, const orbx_keypoint* kps1_un, int n1, const orbx_keypoint* kps2_un, int n2,
                                     const int* matches12, const float* H21, const float* H12, int n_hyp, float sigma,
                                     float* scores, uint8_t* inliers, int* n_list, int* best) {
  if (!c || n_hyp < 1 || !H21 || !H12) return ORBX_ERR_ARG;
  int N = 0;
  int rc = score_stage(c, kps1_un, n1, kps2_un, n2, matches12, &N);
  if (rc) return rc;
  return score_run(c, host_pair_args(c, n_hyp, sigma), 1, false, H21, H12, scores, best, n_list, inliers, N, nullptr, 0);
}

extern "C" int orbx_check_fundamental(orbx_ctx* c, const orbx_keypoint* kps1_un, int n1, const orbx_keypoint* kps2_un, int n2,
                                      const int* matches12, const float* F21, int n_hyp, float sigma, float* scores,
                                      uint8_t* inliers, int* n_list, int* best) {
  if (!c || n_hyp < 1 || !F21) return ORBX_ERR_ARG;
  int N = 0;
  int rc = score_stage(c, kps1_un, n1, kps2_un, n2, matches12, &N);
  if (rc) return rc;
  return score_run(c, host_pair_args(c, n_hyp, sigma), 1, true, F21, nullptr, scores, best, n_list, inliers, N, nullptr, 0);
}

extern "C" int orbx_stream_score_models(orbx_ctx* c, int n_pairs, const float* H21, const float* H12, const float* F21,
                                        int n_hyp, float sigma, float* scores_h, float* scores_f, int* best_h, int* best_f,
                                        uint8_t* inliers_h, uint8_t* inliers_f, int cap, int* n_list) {
  if (!c || n_pairs < 1 || n_hyp < 1 || (H21 && !H12) || (!H21 && !F21)) return ORBX_ERR_ARG;
  if (2 * n_pairs > c->lastBatch) return ORBX_ERR_STATE;  // the pairs must be resident from the last stream run
  if ((inliers_h || inliers_f) && cap < 1) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  const Params& P = c->P;
  ScoreArgs S{};
  S.keysUn = P.keysUn; S.n = P.outN; S.stride = P.kpStride; S.pairStep = 2; S.matches12 = c->dMatches12;
  S.nHyp = n_hyp; S.invSigmaSquare = inv_sigma_square(sigma);
  int rc = ORBX_OK;
  c->stageMs[5] = 0.f;
  if (H21) rc = score_run(c, S, n_pairs, false, H21, H12, scores_h, best_h, n_list, nullptr, 0, inliers_h, cap);
  if (rc) return rc;
  if (F21) rc = score_run(c, S, n_pairs, true, F21, nullptr, scores_f, best_f, n_list, nullptr, 0, inliers_f, cap);
  return rc;
}
