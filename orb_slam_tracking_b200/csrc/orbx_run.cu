// C ABI, part 2: the launch sequences (extract, glue, match), result transfer and inspection.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "ctx.cuh"

using namespace orbx;

// Waits for every step of orbx_stream_process_async still in flight (entry points that touch the slot arrays call it
// first).  A device-side error of a drained step is NOT this call's result: it is kept with the step's ticket, and
// orbx_stream_wait(ticket) reports it -- the caller that drains is usually issuing an unrelated request.
int drain_flights(orbx_ctx* c) {
  for (int t = 0; t < 2; ++t) {
    if (!c->flight[t].active) continue;
    CU(c, cudaEventSynchronize(c->evDone[t]));
    c->flight[t].active = 0;
    int worst = 0;
    for (int i = 0; i < c->flight[t].n; ++i) worst = std::min(worst, c->hStatus[(size_t)t * c->P.maxBatch + i]);
    c->flight[t].status = worst;
  }
  return ORBX_OK;
}

namespace {

enum { ST_PYR = 0, ST_FAST, ST_QUAD, ST_BLUR, ST_DESC, ST_PACK, ST_GLUE, ST_MATCH, ST_N };

inline void mark(orbx_ctx* c, int i) {
  if (c->stageTiming) cudaEventRecord(c->evStage[i], c->stream);
}

// The context's Params with every per-slot array advanced to slot s0, so that a launch over nSlots slots
// works on [s0, s0 + nSlots) (the chunks of orbx_stream_process).
Params shifted(const orbx_ctx* c, int s0) {
  Params Q = c->P;
  if (s0 == 0) return Q;
  const size_t s = (size_t)s0;
  Q.slot0 = s0;
  Q.pyr += s * Q.slabBytes; Q.blur += s * Q.slabBytes;
  Q.candKey += s * Q.candTotal; Q.candXY += s * Q.candTotal; Q.candNode += s * Q.candTotal;
  Q.candCount += s * kMaxLevels;
  if (Q.oldSorted) Q.oldSorted += s * Q.candTotal;
  Q.nodeBox += s * Q.nodeTotal; Q.nodeCount += s * Q.nodeTotal; Q.nodeSlot += s * Q.nodeTotal; Q.nodeBest += s * Q.nodeTotal;
  Q.nodeAux += 4 * s * Q.nodeTotal; Q.nodeList += s * Q.nodeTotal; Q.nodeScan += s * Q.nodeTotal; Q.nodeSort += s * Q.nodeTotal;
  const size_t k = s * (size_t)Q.kpStride;
  Q.lvlXY += k; Q.lvlResp += k; Q.lvlAngle += k; Q.lvlDesc += k * 32; Q.lvlCount += s * kMaxLevels;
  Q.status += s;
  Q.outKp += k; Q.outDesc += k * 32; Q.outN += s; Q.outMono += s;
  Q.keysUn += k; Q.gridOff += s * (kGridCells + 1); Q.gridIdx += k;
  return Q;
}

// Enqueues extraction for the n slots of P (level 0 of each slot must already hold the image).
// beforeOut (optional): an event the stream waits for right before the kernels that overwrite the arrays a previous
// step's download reads (packed keypoints / descriptors, then the frame store and the match lists)
int enqueue_extract(orbx_ctx* c, const Params& P, int n, int lap0, int lap1, cudaStream_t st = nullptr, cudaEvent_t beforeOut = nullptr) {
  if (!st) st = c->stream;
  CU(c, cudaMemsetAsync(P.candCount, 0, (size_t)n * kMaxLevels * sizeof(u32), st));
  CU(c, cudaMemsetAsync(P.status, 0, (size_t)n * sizeof(int), st));
  mark(c, 0);
  c->launches += launch_pyramid(P, c->hTabPtrs, n, st);
  mark(c, 1);
  // the blur only needs the pyramid: it runs on a side stream next to the quadtree (latency-bound, a third of the
  // issue slots; the blur is issue-bound, so they share the SMs well) and joins before the descriptors.  With per-stage timing on,
  // everything stays on one stream so that the stage events bracket one kernel each.
  int side = 0;
  for (int i = 1; i < orbx_ctx::kLanes; ++i) if (st == c->sLane[i]) side = i;
  const bool fork = !c->stageTiming && !P.optSingleStream;
  c->launches += launch_fast(P, c->tma, n, st);
  mark(c, 2);
  if (fork) {  // forked AFTER FAST (itself issue-bound): the blur then shares the SMs with the quadtree only
    CU(c, cudaEventRecord(c->evFork[side], st));
    CU(c, cudaStreamWaitEvent(c->sSide[side], c->evFork[side], 0));
    c->launches += launch_blur(P, n, c->sSide[side]);
    CU(c, cudaEventRecord(c->evJoin[side], c->sSide[side]));
  }
  c->launches += P.kpPolicy == 1 ? launch_retain_old(P, n, st) : launch_quadtree(P, n, st);
  mark(c, 3);
  if (fork) CU(c, cudaStreamWaitEvent(st, c->evJoin[side], 0));
  else c->launches += launch_blur(P, n, st);
  mark(c, 4);
  if (beforeOut) CU(c, cudaStreamWaitEvent(st, beforeOut, 0));
  c->launches += launch_describe(P, c->patchMaps, n, lap0, lap1, st);
  mark(c, 5);
  CU(c, cudaGetLastError());
  c->lastBatch = n;
  return ORBX_OK;
}

// glue + SearchForInitialization over the pairs (2k, 2k+1) of the n slots of P; pair0 = first pair index
int enqueue_glue_match(orbx_ctx* c, const Params& P, int n, int pair0, float nnratio, int checkOri, int window,
                       cudaStream_t st = nullptr) {
  if (!st) st = c->stream;
  c->launches += launch_glue(P, n, st);
  mark(c, 6);
  MatchArgs A;
  A.keysUn = P.keysUn; A.desc = P.outDesc; A.gridOff = P.gridOff; A.gridIdx = P.gridIdx; A.n = P.outN;
  A.stride = P.kpStride;
  A.matches12 = c->dMatches12 + (size_t)pair0 * P.kpStride;
  A.nmatches = c->dNMatches + pair0;
  A.counters = c->dCounters + (size_t)pair0 * 3;
  A.scratch = c->dMatchScratch + (size_t)pair0 * 3 * P.kpStride;
  A.bf = c->dBf;
  A.bf.cand += (size_t)pair0 * P.kpStride; A.bf.candXY += (size_t)pair0 * P.kpStride; A.bf.qIdx += (size_t)pair0 * P.kpStride;
  A.bf.counts += (size_t)pair0 * 4; A.bf.topk += (size_t)pair0 * P.kpStride * 8;
  c->launches += launch_match(P, A, n / 2, 2, nnratio, checkOri, window, st);
  mark(c, 7);
  CU(c, cudaGetLastError());
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

// ---- single-frame latency path -------------------------------------------------------------------------------------
// pinned result block of one frame: n | mono | status | pad | kps[cap] | desc[cap][32] | kps_un[cap] | gridOff[cells + 1] | gridIdx[cap]
struct FrameOut { size_t kps, desc, kun, goff, gidx, total; };
FrameOut frame_out_layout(int cap) {
  FrameOut L;
  size_t o = 16;
  auto take = [&](size_t b) { const size_t r = o; o += (b + 15) & ~(size_t)15; return r; };
  L.kps = take((size_t)cap * sizeof(orbx_keypoint));
  L.desc = take((size_t)cap * 32);
  L.kun = take((size_t)cap * sizeof(orbx_keypoint));
  L.goff = take((kGridCells + 1) * sizeof(u32));
  L.gidx = take((size_t)cap * sizeof(u32));
  L.total = o;
  return L;
}

// Builds (or reuses) the graph of one frame in slot 0.  The stream capture records exactly what the batched path would
// enqueue, so the results are the same bit for bit.
int frame_graph(orbx_ctx* c, int withGlue, int lap0, int lap1, cudaGraphExec_t* out) {
  orbx_ctx::FrameGraph& G = c->fg[withGlue];
  if (G.exec && G.epoch == c->graphEpoch && G.lap0 == lap0 && G.lap1 == lap1) { *out = G.exec; return ORBX_OK; }
  const Params& P = c->P;
  const FrameOut L = frame_out_layout(P.kpStride);
  if (!c->hFrameIn) CU(c, cudaHostAlloc((void**)&c->hFrameIn, (size_t)P.W * P.H, cudaHostAllocDefault));
  if (!c->hFrameOut) CU(c, cudaHostAlloc((void**)&c->hFrameOut, L.total, cudaHostAllocDefault));
  if (G.exec) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
  cudaStream_t st = c->stream;
  CU(c, cudaStreamSynchronize(st));
  const int timing = c->stageTiming;
  c->stageTiming = 0;  // no stage events inside a capture (and the fork to the side stream stays in)
  CU(c, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  int rc = ORBX_OK;
  cudaError_t e = cudaMemcpy2DAsync(P.pyr + P.lv[0].imgOff, P.lv[0].pitch, c->hFrameIn, P.W, P.W, P.H, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) rc = enqueue_extract(c, P, 1, lap0, lap1, st);
  u8* ho = c->hFrameOut;
  const size_t cap = (size_t)P.kpStride;
  if (e == cudaSuccess && rc == ORBX_OK && withGlue) {
    c->launches += launch_glue(P, 1, st);
    e = cudaMemcpyAsync(ho + L.kun, P.keysUn, cap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ho + L.goff, P.gridOff, (kGridCells + 1) * sizeof(u32), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ho + L.gidx, P.gridIdx, cap * sizeof(u32), cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(ho, P.outN, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ho + 4, P.outMono, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ho + 8, P.status, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ho + L.kps, P.outKp, cap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(ho + L.desc, P.outDesc, cap * 32, cudaMemcpyDeviceToHost, st);
  cudaGraph_t graph = nullptr;
  const cudaError_t e2 = cudaStreamEndCapture(st, &graph);
  c->stageTiming = timing;
  if (e != cudaSuccess || e2 != cudaSuccess || rc != ORBX_OK || !graph) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    c->err = std::string("frame graph capture failed: ") + cudaGetErrorString(e != cudaSuccess ? e : e2);
    return rc != ORBX_OK ? rc : ORBX_ERR_CUDA;
  }
  const cudaError_t e3 = cudaGraphInstantiate(&G.exec, graph, 0);
  cudaGraphDestroy(graph);
  if (e3 != cudaSuccess) { G.exec = nullptr; c->err = std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e3); return ORBX_ERR_CUDA; }
  G.lap0 = lap0; G.lap1 = lap1; G.withGlue = withGlue; G.epoch = c->graphEpoch;
  *out = G.exec;
  return ORBX_OK;
}

// One frame through its graph: image -> pinned staging -> ONE graph launch -> ONE synchronisation -> results copied out
int run_frame_graph(orbx_ctx* c, int withGlue, const uint8_t* img, size_t step, const int lapping[2], orbx_keypoint* kps, uint8_t* desc,
                    int cap, int* n_out, int* mono_index, orbx_keypoint* kps_un, uint32_t* grid_offsets, uint32_t* grid_indices) {
  const Params& P = c->P;
  const int lap0 = lapping ? lapping[0] : 0, lap1 = lapping ? lapping[1] : 0;
  int rc = drain_flights(c);
  if (rc) return rc;
  cudaGraphExec_t exec = nullptr;
  if ((rc = frame_graph(c, withGlue, lap0, lap1, &exec))) return rc;
  for (int y = 0; y < P.H; ++y) std::memcpy(c->hFrameIn + (size_t)y * P.W, img + (size_t)y * step, P.W);
  const int launchesPerFrame = c->launches;  // counted during the capture of this graph (or of the other one: same sequence)
  CU(c, cudaGraphLaunch(exec, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  (void)launchesPerFrame;
  const FrameOut L = frame_out_layout(P.kpStride);
  const u8* ho = c->hFrameOut;
  int n, mono, status;
  std::memcpy(&n, ho, 4); std::memcpy(&mono, ho + 4, 4); std::memcpy(&status, ho + 8, 4);
  if (status) { c->err = "device-side status " + std::to_string(status); return status; }
  if (n_out) *n_out = n;
  if (mono_index) *mono_index = mono;
  if (n > cap && (kps || desc || kps_un || grid_indices)) return ORBX_ERR_CAPACITY;
  if (kps && n) std::memcpy(kps, ho + L.kps, (size_t)n * sizeof(orbx_keypoint));
  if (desc && n) std::memcpy(desc, ho + L.desc, (size_t)n * 32);
  if (withGlue) {
    if (kps_un && n) std::memcpy(kps_un, ho + L.kun, (size_t)n * sizeof(orbx_keypoint));
    if (grid_offsets) std::memcpy(grid_offsets, ho + L.goff, (kGridCells + 1) * sizeof(u32));
    if (grid_indices && n) std::memcpy(grid_indices, ho + L.gidx, (size_t)n * sizeof(u32));
  }
  return ORBX_OK;
}

}  // namespace

// Frame::Frame's work in one call (SlamTypes/Frame.cpp:57-77): extraction, undistortion and grid assignment of one frame
// through one CUDA graph and one synchronisation.  Same results as orbx_extract followed by orbx_frame_glue.
extern "C" int orbx_frame_construct(orbx_ctx* c, const uint8_t* img, size_t step, const int lapping[2], orbx_keypoint* kps,
                                    uint8_t* desc, int cap, int* n_out, int* mono_index, orbx_keypoint* kps_un,
                                    uint32_t* grid_offsets, uint32_t* grid_indices) {
  if (!c) return ORBX_ERR_ARG;
  if (!img) return ORBX_ERR_EMPTY_IMAGE;
  if (step < (size_t)c->P.W) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  CU(c, cudaSetDevice(c->device));
  return run_frame_graph(c, 1, img, step, lapping, kps, desc, cap, n_out, mono_index, kps_un, grid_offsets, grid_indices);
}

extern "C" int orbx_extract_batch(orbx_ctx* c, const uint8_t* const* imgs, size_t step, int n_img,
                                  const int lapping[2], orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                                  int* mono_index) {
  if (!c || !imgs || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  const Params& P = c->P;
  for (int i = 0; i < n_img; ++i)
    if (!imgs[i]) return ORBX_ERR_EMPTY_IMAGE;
  if (step < (size_t)P.W) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  // one frame: the latency path (one graph launch, one synchronisation); orbx_set_option("frame_graph", 0) keeps the
  // plain launch sequence (the stage timers need it)
  if (n_img == 1 && P.optFrameGraph && !c->stageTiming)
    return run_frame_graph(c, 0, imgs[0], step, lapping, kps, desc, cap, n_out, mono_index, nullptr, nullptr, nullptr);
  c->launches = 0;
  for (int i = 0; i < n_img; ++i)
    CU(c, cudaMemcpy2DAsync(P.pyr + (size_t)i * P.slabBytes + P.lv[0].imgOff, P.lv[0].pitch, imgs[i], step, P.W, P.H,
                            cudaMemcpyHostToDevice, c->stream));
  const int lap0 = lapping ? lapping[0] : 0, lap1 = lapping ? lapping[1] : 0;
  int rc = enqueue_extract(c, P, n_img, lap0, lap1);
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
  { const int rc = drain_flights(c); if (rc) return rc; }  // ends a chain of asynchronous steps (see process_enqueue)
  // one DMA for the whole batch into the dense staging buffer, then level 0 of every slot on the device
  CU(c, cudaMemcpyAsync(c->dIngest, imgs, (size_t)n_img * P.W * P.H, cudaMemcpyHostToDevice, c->stream));
  launch_ingest(P, c->dIngest, n_img, c->stream);
  CU(c, cudaGetLastError());
  return ORBX_OK;
}

// 3-channel interleaved frames (BGR, or RGB with rgb != 0): Converter::toGray on the device, fused into the ingest
extern "C" int orbx_stream_upload_color(orbx_ctx* c, const uint8_t* imgs, int n_img, int rgb) {
  if (!c || !imgs || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  const Params& P = c->P;
  CU(c, cudaSetDevice(c->device));
  { const int rc = drain_flights(c); if (rc) return rc; }
  if (!c->dIngestColor) {
    void* q = nullptr;
    CU(c, cudaMalloc(&q, (size_t)P.maxBatch * P.W * P.H * 3));
    c->allocs.push_back(q);
    c->dIngestColor = (u8*)q;
  }
  CU(c, cudaMemcpyAsync(c->dIngestColor, imgs, (size_t)n_img * P.W * P.H * 3, cudaMemcpyHostToDevice, c->stream));
  launch_ingest_color(P, c->dIngestColor, rgb, n_img, c->stream);
  CU(c, cudaGetLastError());
  return ORBX_OK;
}

extern "C" int orbx_stream_run(orbx_ctx* c, int n_img, float nnratio, int check_orientation, int window_size) {
  if (!c || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  const Params& P = c->P;
  CU(c, cudaSetDevice(c->device));
  c->launches = 0;
  int rc;
  if ((rc = drain_flights(c))) return rc;
  const int parts = std::min(orbx_ctx::kLanes, std::max(1, P.optRunParts > 0 ? P.optRunParts : 2));
  if (n_img >= 64 * parts && parts > 1 && !c->stageTiming && !P.optSingleStream) {
    // the batch in `parts` pieces (each a whole number of pairs), one per compute lane: the latency-bound kernels of
    // one piece (quadtree, matcher) overlap the issue-bound ones of the others; all pieces fence on the context's stream
    CU(c, cudaEventRecord(c->evLane[0], c->stream));
    int s0 = 0;
    for (int k = 0; k < parts; ++k) {
      const int s1 = k == parts - 1 ? n_img : (int)((long)n_img * (k + 1) / parts) & ~1;
      cudaStream_t st = c->sLane[k];
      if (k) CU(c, cudaStreamWaitEvent(st, c->evLane[0], 0));
      const Params Q = shifted(c, s0);
      if ((rc = enqueue_extract(c, Q, s1 - s0, 0, 0, st))) return rc;
      if ((rc = enqueue_glue_match(c, Q, s1 - s0, s0 / 2, nnratio, check_orientation, window_size, st))) return rc;
      if (k) {
        CU(c, cudaEventRecord(c->evLane[k], st));
        CU(c, cudaStreamWaitEvent(c->stream, c->evLane[k], 0));
      }
      s0 = s1;
    }
    c->lastBatch = n_img;
    return ORBX_OK;
  }
  if ((rc = enqueue_extract(c, P, n_img, 0, 0))) return rc;
  return enqueue_glue_match(c, P, n_img, 0, nnratio, check_orientation, window_size);
}

extern "C" int orbx_stream_sync(orbx_ctx* c) {
  if (!c) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  const int flights = drain_flights(c);
  if (flights) return flights;
  CU(c, cudaStreamSynchronize(c->stream));
  if (c->stageTiming) {
    static const int a[ST_N] = {0, 1, 2, 3, 4, 5, 5, 6}, b[ST_N] = {1, 2, 3, 4, 5, 5, 6, 7};
    for (int i = 0; i < ST_N; ++i) {
      float ms = 0.f;
      if (a[i] == b[i]) continue;  // slot 5 belongs to orbx_stream_score_models
      cudaEventElapsedTime(&ms, c->evStage[a[i]], c->evStage[b[i]]);
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
  // rows are copied whole: a row shorter than a frame's keypoint capacity would truncate silently (and matches12 could then
  // hold indices beyond it), so such a buffer is refused up front
  if ((kps || kps_un || desc || matches12) && cap < P.kpStride) return ORBX_ERR_CAPACITY;
  const int m = std::min(cap, P.kpStride);
  const size_t w = (size_t)m;
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

// Host images in, results out: the batch is cut into chunks of `chunk` frames and the three legs of each chunk run on
// their own streams -- H2D DMA of chunk k+1 and D2H DMA of chunk k-1 overlap the kernels of chunk k.  imgs and the
// output buffers should be page-locked (orbx_host_alloc) for the copies to be asynchronous.
//
// process_enqueue only enqueues; up to two steps may be in flight (orbx_stream_process_async / orbx_stream_wait), so that
// the upload of step s+1 overlaps the kernels and the download of step s.  What keeps two steps apart on the device:
//   * chunk j of every step runs on compute lane j % lanes: stream order separates its kernels from the previous step's;
//   * the H2D stream waits for evIngested[j] (the previous step's k_ingest has consumed the staging area of chunk j);
//   * the lane waits for evStatus[j] before it clears the status words, and for evD2H[j] (the previous step's results of
//     chunk j have left the device) right before the kernels that overwrite them (k_pack, k_glue, k_match).
// A step with another chunk schedule than the one in flight drains the pipeline first.
namespace {


int process_enqueue(orbx_ctx* c, const uint8_t* imgs, int n_img, float nnratio, int check_orientation, int window_size,
                    int chunk, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc, int cap, int* n_out, int* matches12,
                    int* nmatches, int* counters3, int ticket, bool streaming) {
  const Params& P = c->P;
  // default chunking: a blocking call 128 frames (ramped, below); a stream of steps three equal chunks, one per lane
  // (measured at 512 x 640x480: 3 x 172 on 3 lanes 145.7 k frames/s, 2 x 256 on 2 lanes 145.0 k, 4 x 128 on 2 lanes 133.4 k)
  if (chunk <= 0) chunk = streaming ? ((n_img + 2) / 3 + 1) & ~1 : 128;
  chunk = std::max(2, chunk & ~1);
  // chunk boundaries.  A blocking call starts with short chunks (chunk / 4, then chunk / 2) so that the kernels start
  // early; in a stream of steps the ramp is hidden behind the previous step and equal chunks run the kernels better
  std::vector<int> cut(1, 0);
  for (int sz = streaming ? chunk : std::max(2, (chunk / 4) & ~1); cut.back() < n_img; sz = std::min(chunk, 2 * sz)) cut.push_back(std::min(n_img, cut.back() + sz));
  const int nChunks = (int)cut.size() - 1;
  if (c->flight[ticket ^ 1].active && (c->flight[ticket ^ 1].n != n_img || c->flight[ticket ^ 1].chunk != chunk)) {
    const int rc = drain_flights(c);
    if (rc) return rc;
  }
  while ((int)c->evChunk.size() < 5 * nChunks) {
    cudaEvent_t e;
    CU(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c->evChunk.push_back(e);
  }
  const bool chained = c->flight[ticket ^ 1].active != 0;  // the other step is still in flight with the same schedule
  const size_t imgBytes = (size_t)P.W * P.H;
  const int m = std::min(cap, P.kpStride);
  const size_t w = (size_t)m;
  const int timing = c->stageTiming;
  c->stageTiming = 0;  // per-stage events belong to the unchunked orbx_stream_run
  c->launches = 0;
  int* hStatus = c->hStatus + (size_t)ticket * P.maxBatch;
  const int lanesEnv = P.optLanes;
  // A stream of steps stays off lane 0: lane 0 is the context's stream, which joins every step's fence (below) so that
  // later calls on it come after the step -- as a compute lane it would stall the next step behind this step's download.
  const int lane0 = streaming ? 1 : 0;
  const int lanes = std::min(orbx_ctx::kLanes - lane0, std::max(1, lanesEnv > 0 ? lanesEnv : 3));
  int rc = ORBX_OK;
  if (!chained) {  // whatever the caller enqueued on the context's stream (orbx_stream_run ...) comes first.  A chained step needs no
                   // fence: every other entry point that enqueues on the context's stream drains the steps in flight first, so
                   // nothing can sit between two chained steps
    CU(c, cudaEventRecord(c->evLane[0], c->stream));
    for (int i = 0; i < lanes; ++i)
      if (lane0 + i) CU(c, cudaStreamWaitEvent(c->sLane[lane0 + i], c->evLane[0], 0));
    CU(c, cudaStreamWaitEvent(c->sH2D, c->evLane[0], 0));
  }
  for (int k = 0; k < nChunks && rc == ORBX_OK; ++k) {
    const int s0 = cut[k], n = cut[k + 1] - s0, p0 = s0 / 2, np = n / 2;
    cudaEvent_t evIn = c->evChunk[5 * k], evOut = c->evChunk[5 * k + 1], evIngested = c->evChunk[5 * k + 2], evD2H = c->evChunk[5 * k + 3],
                evStatus = c->evChunk[5 * k + 4];
    if (chained) CU(c, cudaStreamWaitEvent(c->sH2D, evIngested, 0));
    CU(c, cudaMemcpyAsync(c->dIngest + s0 * imgBytes, imgs + s0 * imgBytes, n * imgBytes, cudaMemcpyHostToDevice, c->sH2D));
    CU(c, cudaEventRecord(evIn, c->sH2D));
    // consecutive chunks rotate over the compute lanes: the latency-bound tail of a chunk (quadtree, matcher: few
    // CTAs) overlaps the issue-bound head of the next ones; chunks touch disjoint slots
    cudaStream_t sc = c->sLane[lane0 + k % lanes];
    CU(c, cudaStreamWaitEvent(sc, evIn, 0));
    if (chained) CU(c, cudaStreamWaitEvent(sc, evStatus, 0));  // the previous step's status words of these slots are on the host
    const Params Q = shifted(c, s0);
    c->launches += launch_ingest(Q, c->dIngest + s0 * imgBytes, n, sc);
    CU(c, cudaEventRecord(evIngested, sc));
    // the previous step's results of these slots must have left the device before k_pack / k_glue / k_match overwrite them
    if ((rc = enqueue_extract(c, Q, n, 0, 0, sc, chained ? evD2H : nullptr))) break;
    if ((rc = enqueue_glue_match(c, Q, n, p0, nnratio, check_orientation, window_size, sc))) break;
    CU(c, cudaEventRecord(evOut, sc));
    cudaStream_t sd = c->sD2H;
    CU(c, cudaStreamWaitEvent(sd, evOut, 0));
    CU(c, cudaMemcpyAsync(hStatus + s0, Q.status, n * sizeof(int), cudaMemcpyDeviceToHost, sd));
    CU(c, cudaEventRecord(evStatus, sd));
    if (kps)
      CU(c, cudaMemcpy2DAsync(kps + (size_t)s0 * cap, (size_t)cap * sizeof(orbx_keypoint), Q.outKp,
                              (size_t)P.kpStride * sizeof(orbx_keypoint), w * sizeof(orbx_keypoint), n, cudaMemcpyDeviceToHost, sd));
    if (kps_un)
      CU(c, cudaMemcpy2DAsync(kps_un + (size_t)s0 * cap, (size_t)cap * sizeof(orbx_keypoint), Q.keysUn,
                              (size_t)P.kpStride * sizeof(orbx_keypoint), w * sizeof(orbx_keypoint), n, cudaMemcpyDeviceToHost, sd));
    if (desc)
      CU(c, cudaMemcpy2DAsync(desc + (size_t)s0 * cap * 32, (size_t)cap * 32, Q.outDesc, (size_t)P.kpStride * 32, w * 32, n,
                              cudaMemcpyDeviceToHost, sd));
    if (n_out) CU(c, cudaMemcpyAsync(n_out + s0, Q.outN, n * sizeof(int), cudaMemcpyDeviceToHost, sd));
    if (np > 0) {
      if (matches12)
        CU(c, cudaMemcpy2DAsync(matches12 + (size_t)p0 * cap, (size_t)cap * sizeof(int), c->dMatches12 + (size_t)p0 * P.kpStride,
                                (size_t)P.kpStride * sizeof(int), w * sizeof(int), np, cudaMemcpyDeviceToHost, sd));
      if (nmatches) CU(c, cudaMemcpyAsync(nmatches + p0, c->dNMatches + p0, np * sizeof(int), cudaMemcpyDeviceToHost, sd));
      if (counters3) CU(c, cudaMemcpyAsync(counters3 + 3 * p0, c->dCounters + 3 * p0, np * 3 * sizeof(int), cudaMemcpyDeviceToHost, sd));
    }
    CU(c, cudaEventRecord(evD2H, sd));
  }
  c->stageTiming = timing;
  // the D2H stream is serial and each of its copies waits for its chunk's kernels: its last event fences the whole step.
  // The context's stream (lane 0) joins it, so work enqueued on it later -- and orbx_timer_stop -- comes after the step.
  CU(c, cudaEventRecord(c->evDone[ticket], c->sD2H));
  CU(c, cudaStreamWaitEvent(c->stream, c->evDone[ticket], 0));
  c->flight[ticket].active = 1; c->flight[ticket].n = n_img; c->flight[ticket].chunk = chunk; c->flight[ticket].status = 0;
  if (rc) { drain_flights(c); return rc; }
  c->lastBatch = n_img;
  return ORBX_OK;
}

int wait_ticket(orbx_ctx* c, int ticket) {
  if (!c->flight[ticket].active) {  // already drained by another call: its status was kept for this one
    const int st = c->flight[ticket].status;
    c->flight[ticket].status = 0;
    if (st) c->err = "device-side status " + std::to_string(st);
    return st;
  }
  CU(c, cudaEventSynchronize(c->evDone[ticket]));
  c->flight[ticket].active = 0;
  CU(c, cudaGetLastError());
  int worst = 0;
  const int* hStatus = c->hStatus + (size_t)ticket * c->P.maxBatch;
  for (int i = 0; i < c->flight[ticket].n; ++i) worst = std::min(worst, hStatus[i]);
  if (worst) c->err = "device-side status " + std::to_string(worst);
  return worst;
}

}  // namespace

extern "C" int orbx_stream_process(orbx_ctx* c, const uint8_t* imgs, int n_img, float nnratio, int check_orientation,
                                   int window_size, int chunk, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc,
                                   int cap, int* n_out, int* matches12, int* nmatches, int* counters3) {
  if (!c || !imgs || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  if ((kps || kps_un || desc || matches12) && cap < c->P.kpStride) return ORBX_ERR_CAPACITY;  // see orbx_stream_download
  CU(c, cudaSetDevice(c->device));
  int rc = drain_flights(c);
  if (rc) return rc;
  rc = process_enqueue(c, imgs, n_img, nnratio, check_orientation, window_size, chunk, kps, kps_un, desc, cap, n_out, matches12,
                       nmatches, counters3, 0, false);
  if (rc) return rc;
  return wait_ticket(c, 0);
}

// Streaming form: returns a ticket (0 or 1) as soon as the step is enqueued; at most two steps are in flight, and the
// caller owns two sets of output buffers.  orbx_stream_wait(ticket) blocks until that step's results are on the host.
extern "C" int orbx_stream_process_async(orbx_ctx* c, const uint8_t* imgs, int n_img, float nnratio, int check_orientation,
                                         int window_size, int chunk, orbx_keypoint* kps, orbx_keypoint* kps_un, uint8_t* desc,
                                         int cap, int* n_out, int* matches12, int* nmatches, int* counters3) {
  if (!c || !imgs || n_img < 1 || n_img > c->P.maxBatch) return ORBX_ERR_ARG;
  if (!c->cameraSet) return ORBX_ERR_STATE;
  if ((kps || kps_un || desc || matches12) && cap < c->P.kpStride) return ORBX_ERR_CAPACITY;
  CU(c, cudaSetDevice(c->device));
  const int ticket = c->nextTicket;
  if (c->flight[ticket].active) return ORBX_ERR_STATE;  // wait for this ticket first
  const int rc = process_enqueue(c, imgs, n_img, nnratio, check_orientation, window_size, chunk, kps, kps_un, desc, cap, n_out,
                                 matches12, nmatches, counters3, ticket, true);
  if (rc) return rc;
  c->nextTicket = ticket ^ 1;
  return ticket;
}

extern "C" int orbx_stream_wait(orbx_ctx* c, int ticket) {
  if (!c || ticket < 0 || ticket > 1) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  return wait_ticket(c, ticket);
}

// ---- timing ---------------------------------------------------------------------------------------
extern "C" int orbx_timer_start(orbx_ctx* c) {
  if (!c) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  CU(c, cudaEventRecord(c->evStart, c->stream));
  return ORBX_OK;
}
extern "C" int orbx_timer_stop(orbx_ctx* c, float* ms) {
  if (!c || !ms) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
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

// Copy-only probe of the host link (no kernels): `reps` times the uploads and downloads of one e2e step -- in_bytes from
// pinned host memory to the device on the H2D stream, out_bytes back on the D2H stream, both directions in flight at
// once -- timed on the device.  With every rank of a multi-GPU job calling it at the same time it measures the link the
// ranks share; bench.py reports the e2e step time against it (e2e.host_link_frac).
extern "C" int orbx_hostlink_probe(orbx_ctx* c, const void* host_in, size_t in_bytes, void* host_out, size_t out_bytes, int reps,
                                   float* ms_per_rep) {
  if (!c || !host_in || !host_out || reps < 1 || !ms_per_rep) return ORBX_ERR_ARG;
  const size_t cap = (size_t)c->P.maxBatch * c->P.W * c->P.H;
  if (in_bytes > cap || out_bytes > (size_t)c->P.maxBatch * c->P.slabBytes) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  int rc = drain_flights(c);
  if (rc) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  cudaEvent_t e0, e1, eUp, eDn;
  CU(c, cudaEventCreate(&e0)); CU(c, cudaEventCreate(&e1));
  CU(c, cudaEventCreateWithFlags(&eUp, cudaEventDisableTiming)); CU(c, cudaEventCreateWithFlags(&eDn, cudaEventDisableTiming));
  CU(c, cudaEventRecord(e0, c->stream));
  CU(c, cudaStreamWaitEvent(c->sH2D, e0, 0));
  CU(c, cudaStreamWaitEvent(c->sD2H, e0, 0));
  for (int i = 0; i < reps; ++i) {
    CU(c, cudaMemcpyAsync(c->dIngest, host_in, in_bytes, cudaMemcpyHostToDevice, c->sH2D));
    CU(c, cudaMemcpyAsync(host_out, c->P.blur, out_bytes, cudaMemcpyDeviceToHost, c->sD2H));  // any device bytes: only the link is measured
  }
  CU(c, cudaEventRecord(eUp, c->sH2D));
  CU(c, cudaEventRecord(eDn, c->sD2H));
  CU(c, cudaStreamWaitEvent(c->stream, eUp, 0));
  CU(c, cudaStreamWaitEvent(c->stream, eDn, 0));
  CU(c, cudaEventRecord(e1, c->stream));
  CU(c, cudaEventSynchronize(e1));
  float ms = 0.f;
  CU(c, cudaEventElapsedTime(&ms, e0, e1));
  *ms_per_rep = ms / reps;
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(eUp); cudaEventDestroy(eDn);
  return ORBX_OK;
}

// pinned host memory for the e2e path
extern "C" void* orbx_host_alloc(size_t bytes) {
  void* p = nullptr;
  return cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? p : nullptr;
}
// write_combined != 0: cudaHostAllocWriteCombined (the CPU only writes the buffer, e.g. frames on their way to the device:
// the DMA reads need no cache snoop); such a buffer is slow to READ from the CPU
extern "C" void* orbx_host_alloc_ex(size_t bytes, int write_combined) {
  void* p = nullptr;
  return cudaHostAlloc(&p, bytes ? bytes : 1, write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault) == cudaSuccess ? p : nullptr;
}
extern "C" void orbx_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
