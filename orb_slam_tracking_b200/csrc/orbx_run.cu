// C ABI, part 2: the launch sequences (extract, glue, match), result transfer and inspection.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "ctx.cuh"

using namespace orbx;

namespace {

enum { ST_PYR = 0, ST_FAST, ST_QUAD, ST_BLUR, ST_DESC, ST_PACK, ST_GLUE, ST_MATCH, ST_N };

inline void mark(orbx_ctx* c, int i) {
  if (c->stageTiming) cudaEventRecord(c->evStage[i], c->stream);
}

// Enqueues extraction for slots [0, n).  Level 0 of each slot must already hold the image.
int enqueue_extract(orbx_ctx* c, int n, int lap0, int lap1) {
  const Params& P = c->P;
  cudaStream_t st = c->stream;
  CU(c, cudaMemsetAsync(P.candCount, 0, (size_t)n * kMaxLevels * sizeof(u32), st));
  CU(c, cudaMemsetAsync(P.status, 0, (size_t)n * sizeof(int), st));
  mark(c, 0);
  c->launches += launch_pyramid(P, c->hTabPtrs, n, st);
  mark(c, 1);
  c->launches += launch_fast(P, c->tma, n, st);
  mark(c, 2);
  c->launches += launch_quadtree(P, n, st);
  mark(c, 3);
  c->launches += launch_blur(P, n, st);
  mark(c, 4);
  c->launches += launch_describe(P, n, lap0, lap1, st);
  mark(c, 5);
  CU(c, cudaGetLastError());
  c->lastBatch = n;
  return ORBX_OK;
}

int check_status(orbx_ctx* c, int n) {
  std::vector<int> st(n);
  CU(c, cudaMemcpyAsync(st.data(), c->P.status, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  int worst = 0;
  for (int i = 0; i < n; ++i) worst = std::min(worst, st[i]);
  if (worst) c->err = "device-side status " + std::to_string(worst);
  return worst;
}

}  // namespace

extern "C" int orbx_extract_batch(orbx_ctx* c, const uint8_t* const* imgs, size_t step, int n_img,
                                  const int lapping[2], orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                                  int* mono_index) {
  if (!c || !imgs || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  const Params& P = c->P;
  for (int i = 0; i < n_img; ++i)
    if (!imgs[i]) return ORBX_ERR_EMPTY_IMAGE;
  if (step < (size_t)P.W) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  c->launches = 0;
  for (int i = 0; i < n_img; ++i)
    CU(c, cudaMemcpy2DAsync(P.pyr + (size_t)i * P.slabBytes + P.lv[0].imgOff, P.lv[0].pitch, imgs[i], step, P.W, P.H,
                            cudaMemcpyHostToDevice, c->stream));
  const int lap0 = lapping ? lapping[0] : 0, lap1 = lapping ? lapping[1] : 0;
  int rc = enqueue_extract(c, n_img, lap0, lap1);
  if (rc) return rc;
  std::vector<int> n(n_img), mono(n_img);
  CU(c, cudaMemcpyAsync(n.data(), P.outN, n_img * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(mono.data(), P.outMono, n_img * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  if ((rc = check_status(c, n_img))) return rc;
  for (int i = 0; i < n_img; ++i) {
    if (n_out) n_out[i] = n[i];
    if (mono_index) mono_index[i] = mono[i];
    if (n[i] > cap && (kps || desc)) return ORBX_ERR_CAPACITY;
    if (kps && n[i])
      CU(c, cudaMemcpyAsync(kps + (size_t)i * cap, P.outKp + (size_t)i * P.kpStride, (size_t)n[i] * sizeof(orbx_keypoint),
                            cudaMemcpyDeviceToHost, c->stream));
    if (desc && n[i])
      CU(c, cudaMemcpyAsync(desc + (size_t)i * cap * 32, P.outDesc + (size_t)i * P.kpStride * 32, (size_t)n[i] * 32,
                            cudaMemcpyDeviceToHost, c->stream));
  }
  CU(c, cudaStreamSynchronize(c->stream));
  return ORBX_OK;
}

extern "C" int orbx_extract(orbx_ctx* c, const uint8_t* img, size_t step, const int lapping[2], orbx_keypoint* kps,
                            uint8_t* desc, int cap, int* n_out, int* mono_index) {
  if (!img) return ORBX_ERR_EMPTY_IMAGE;
  const uint8_t* one[1] = {img};
  return orbx_extract_batch(c, one, step, 1, lapping, kps, desc, cap, n_out, mono_index);
}

// ---- inspection ---------------------------------------------------------------------------------
static int get_level(orbx_ctx* c, const u8* base, int slot, int level, int padded, uint8_t* dst, size_t dst_step) {
  if (!c || !dst || slot < 0 || slot >= c->P.maxBatch || level < 0 || level >= c->P.nlevels) return ORBX_ERR_ARG;
  const LevelInfo& L = c->P.lv[level];
  const int b = padded ? kEdge : 0;
  if (dst_step < (size_t)(L.w + 2 * b)) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->stream));
  uint8_t* roi = dst + (size_t)b * dst_step + b;
  CU(c, cudaMemcpy2D(roi, dst_step, base + (size_t)slot * c->P.slabBytes + L.imgOff, L.pitch, L.w, L.h,
                     cudaMemcpyDeviceToHost));
  if (padded) {
    // cv::copyMakeBorder(..., 19, BORDER_REFLECT_101) of the level itself (:1689-1708); a pure copy
    // of pixels already produced on the device, done while handing the level to the host.
    for (int y = 0; y < L.h; ++y) {
      uint8_t* row = roi + (size_t)y * dst_step;
      for (int k = 1; k <= b; ++k) { row[-k] = row[k]; row[L.w - 1 + k] = row[L.w - 1 - k]; }
    }
    for (int k = 1; k <= b; ++k) {
      std::memcpy(roi - b + (ptrdiff_t)(-k) * (ptrdiff_t)dst_step, roi - b + (size_t)k * dst_step, L.w + 2 * b);
      std::memcpy(roi - b + (size_t)(L.h - 1 + k) * dst_step, roi - b + (size_t)(L.h - 1 - k) * dst_step, L.w + 2 * b);
    }
  }
  return ORBX_OK;
}
extern "C" int orbx_get_pyramid_level(orbx_ctx* c, int slot, int level, int padded, uint8_t* dst, size_t dst_step) {
  return c ? get_level(c, c->P.pyr, slot, level, padded, dst, dst_step) : ORBX_ERR_ARG;
}
extern "C" int orbx_get_blurred_level(orbx_ctx* c, int slot, int level, uint8_t* dst, size_t dst_step) {
  return c ? get_level(c, c->P.blur, slot, level, 0, dst, dst_step) : ORBX_ERR_ARG;
}

extern "C" int orbx_get_candidates(orbx_ctx* c, int slot, int level, float* xyr, int cap) {
  if (!c || slot < 0 || slot >= c->P.maxBatch || level < 0 || level >= c->P.nlevels) return ORBX_ERR_ARG;
  const Params& P = c->P;
  const LevelInfo& L = P.lv[level];
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->stream));
  u32 n = 0;
  CU(c, cudaMemcpy(&n, P.candCount + (size_t)slot * kMaxLevels + level, sizeof(u32), cudaMemcpyDeviceToHost));
  n = std::min(n, (u32)L.candCap);
  std::vector<u64> key(n);
  std::vector<u32> xy(n);
  const size_t off = (size_t)slot * P.candTotal + L.candOff;
  if (n) {
    CU(c, cudaMemcpy(key.data(), P.candKey + off, n * sizeof(u64), cudaMemcpyDeviceToHost));
    CU(c, cudaMemcpy(xy.data(), P.candXY + off, n * sizeof(u32), cudaMemcpyDeviceToHost));
  }
  // the device list is unordered; the order field of the key restores the reference's emission order
  std::vector<u32> idx(n);
  for (u32 i = 0; i < n; ++i) idx[i] = i;
  std::sort(idx.begin(), idx.end(), [&](u32 a, u32 b) { return (u32)~key[a] < (u32)~key[b]; });
  for (u32 i = 0; i < n && (int)i < cap; ++i) {
    const u32 j = idx[i];
    xyr[3 * i] = (float)(xy[j] & 0xffff);
    xyr[3 * i + 1] = (float)(xy[j] >> 16);
    xyr[3 * i + 2] = (float)(u32)(key[j] >> 32);
  }
  return (int)n;
}

extern "C" int orbx_get_level_keypoints(orbx_ctx* c, int slot, int level, orbx_keypoint* kps, uint8_t* desc, int cap) {
  if (!c || slot < 0 || slot >= c->P.maxBatch || level < 0 || level >= c->P.nlevels) return ORBX_ERR_ARG;
  const Params& P = c->P;
  const LevelInfo& L = P.lv[level];
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->stream));
  int n = 0;
  CU(c, cudaMemcpy(&n, P.lvlCount + (size_t)slot * kMaxLevels + level, sizeof(int), cudaMemcpyDeviceToHost));
  const int m = std::min(n, cap);
  if (m <= 0) return n;
  const size_t off = (size_t)slot * P.kpStride + L.kpOff;
  std::vector<u32> xy(m);
  std::vector<float> resp(m), ang(m);
  CU(c, cudaMemcpy(xy.data(), P.lvlXY + off, m * sizeof(u32), cudaMemcpyDeviceToHost));
  CU(c, cudaMemcpy(resp.data(), P.lvlResp + off, m * sizeof(float), cudaMemcpyDeviceToHost));
  CU(c, cudaMemcpy(ang.data(), P.lvlAngle + off, m * sizeof(float), cudaMemcpyDeviceToHost));
  if (kps)
    for (int i = 0; i < m; ++i) {
      orbx_keypoint k;
      k.x = (float)(xy[i] & 0xffff); k.y = (float)(xy[i] >> 16); k.size = L.patchSize; k.angle = ang[i];
      k.response = resp[i]; k.octave = level; k.class_id = -1;
      kps[i] = k;
    }
  if (desc) CU(c, cudaMemcpy(desc, P.lvlDesc + off * 32, (size_t)m * 32, cudaMemcpyDeviceToHost));
  return n;
}

// ---- device-resident stream path ---------------------------------------------------------------
extern "C" int orbx_stream_upload(orbx_ctx* c, const uint8_t* imgs, int n_img) {
  if (!c || !imgs || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  const Params& P = c->P;
  CU(c, cudaSetDevice(c->device));
  // one strided copy for the whole batch: rows of W bytes, W*H apart per image -> pitch rows, slab apart
  for (int i = 0; i < n_img; ++i)
    CU(c, cudaMemcpy2DAsync(P.pyr + (size_t)i * P.slabBytes + P.lv[0].imgOff, P.lv[0].pitch,
                            imgs + (size_t)i * P.W * P.H, P.W, P.W, P.H, cudaMemcpyHostToDevice, c->stream));
  return ORBX_OK;
}

extern "C" int orbx_stream_run(orbx_ctx* c, int n_img, float nnratio, int check_orientation, int window_size) {
  if (!c || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  const Params& P = c->P;
  CU(c, cudaSetDevice(c->device));
  c->launches = 0;
  int rc = enqueue_extract(c, n_img, 0, 0);
  if (rc) return rc;
  c->launches += launch_glue(P, n_img, c->stream);
  mark(c, 6);
  MatchArgs A;
  A.keysUn = P.keysUn; A.desc = P.outDesc; A.gridOff = P.gridOff; A.gridIdx = P.gridIdx; A.n = P.outN;
  A.stride = P.kpStride; A.matches12 = c->dMatches12; A.nmatches = c->dNMatches; A.counters = c->dCounters;
  A.scratch = c->dMatchScratch;
  c->launches += launch_match(P, A, n_img / 2, 2, nnratio, check_orientation, window_size, c->stream);
  mark(c, 7);
  CU(c, cudaGetLastError());
  return ORBX_OK;
}

extern "C" int orbx_stream_sync(orbx_ctx* c) {
  if (!c) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaStreamSynchronize(c->stream));
  if (c->stageTiming) {
    static const int a[ST_N] = {0, 1, 2, 3, 4, 5, 5, 6}, b[ST_N] = {1, 2, 3, 4, 5, 5, 6, 7};
    for (int i = 0; i < ST_N; ++i) {
      float ms = 0.f;
      if (a[i] != b[i]) cudaEventElapsedTime(&ms, c->evStage[a[i]], c->evStage[b[i]]);
      c->stageMs[i] = ms;
    }
    cudaGetLastError();
  }
  return ORBX_OK;
}

extern "C" int orbx_stream_download(orbx_ctx* c, int n_img, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc,
                                    int cap, int* n_out, int* matches12, int* nmatches, int* counters3) {
  if (!c || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  const Params& P = c->P;
  CU(c, cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  int rc = check_status(c, n_img);
  if (rc) return rc;
  const int m = std::min(cap, P.kpStride);
  const size_t w = (size_t)m;
  if ((kps || kps_un || desc || matches12) && cap < 1) return ORBX_ERR_ARG;
  if (kps)
    CU(c, cudaMemcpy2DAsync(kps, (size_t)cap * sizeof(orbx_keypoint), P.outKp, (size_t)P.kpStride * sizeof(orbx_keypoint),
                            w * sizeof(orbx_keypoint), n_img, cudaMemcpyDeviceToHost, st));
  if (kps_un)
    CU(c, cudaMemcpy2DAsync(kps_un, (size_t)cap * sizeof(orbx_keypoint), P.keysUn, (size_t)P.kpStride * sizeof(orbx_keypoint),
                            w * sizeof(orbx_keypoint), n_img, cudaMemcpyDeviceToHost, st));
  if (desc)
    CU(c, cudaMemcpy2DAsync(desc, (size_t)cap * 32, P.outDesc, (size_t)P.kpStride * 32, w * 32, n_img,
                            cudaMemcpyDeviceToHost, st));
  if (n_out) CU(c, cudaMemcpyAsync(n_out, P.outN, n_img * sizeof(int), cudaMemcpyDeviceToHost, st));
  const int nPairs = n_img / 2;
  if (nPairs > 0) {
    if (matches12)
      CU(c, cudaMemcpy2DAsync(matches12, (size_t)cap * sizeof(int), c->dMatches12, (size_t)P.kpStride * sizeof(int),
                              w * sizeof(int), nPairs, cudaMemcpyDeviceToHost, st));
    if (nmatches) CU(c, cudaMemcpyAsync(nmatches, c->dNMatches, nPairs * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (counters3) CU(c, cudaMemcpyAsync(counters3, c->dCounters, nPairs * 3 * sizeof(int), cudaMemcpyDeviceToHost, st));
  }
  CU(c, cudaStreamSynchronize(st));
  return ORBX_OK;
}

// ---- timing ---------------------------------------------------------------------------------------
extern "C" int orbx_timer_start(orbx_ctx* c) {
  if (!c) return ORBX_ERR_ARG;
  CU(c, cudaEventRecord(c->evStart, c->stream));
  return ORBX_OK;
}
extern "C" int orbx_timer_stop(orbx_ctx* c, float* ms) {
  if (!c || !ms) return ORBX_ERR_ARG;
  CU(c, cudaEventRecord(c->evStop, c->stream));
  CU(c, cudaEventSynchronize(c->evStop));
  CU(c, cudaEventElapsedTime(ms, c->evStart, c->evStop));
  return ORBX_OK;
}
extern "C" int orbx_set_stage_timing(orbx_ctx* c, int enable) {
  if (!c) return ORBX_ERR_ARG;
  c->stageTiming = enable != 0;
  return ORBX_OK;
}
extern "C" int orbx_get_stage_times(orbx_ctx* c, float* ms8) {
  if (!c || !ms8) return ORBX_ERR_ARG;
  std::memcpy(ms8, c->stageMs, sizeof(c->stageMs));
  return ORBX_OK;
}

// pinned host memory for the e2e path
extern "C" void* orbx_host_alloc(size_t bytes) {
  void* p = nullptr;
  return cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? p : nullptr;
}
extern "C" void orbx_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
