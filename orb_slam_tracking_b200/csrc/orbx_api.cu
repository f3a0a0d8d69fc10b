// C ABI of liborbx.so (include/orbx.h): context lifetime, host<->device plumbing, launch sequence.
// No algorithm lives here; each stage is a kernel in pyramid/fast/quadtree/blur/describe/glue/match.cu.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "ctx.cuh"

namespace orbx {
void build_resize_tab(int sw, int sh, int dw, int dh, std::vector<int>& tab);
int build_geometry(Params& P, const HostTables& T, int fastTileBudget, std::vector<StripDesc>& strips, std::vector<BlurTile>& tiles,
                   std::vector<u8>& rootTab, std::vector<int>& rootEdges);
int configure_fast(Params& P);
int build_strips_old(Params& P, const HostTables& T, int fastTileBudget, std::vector<StripDesc>& strips, int& tileMaxOut, int& segWOut,
                     int& segPOut, int& cellTotal);
}
using namespace orbx;

template <typename T>
static int dalloc(orbx_ctx* c, T** p, size_t n) {
  void* q = nullptr;
  if (n == 0) n = 1;
  CU(c, cudaMalloc(&q, n * sizeof(T)));
  c->allocs.push_back(q);
  *p = (T*)q;
  return ORBX_OK;
}
template <typename T>
static int dupload(orbx_ctx* c, T** p, const std::vector<T>& v) {
  int rc = dalloc(c, p, v.size());
  if (rc) return rc;
  if (!v.empty()) CU(c, cudaMemcpy(*p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return ORBX_OK;
}

// One u8 tensor map per pyramid level, {x, y, slot}; the driver entry point is taken from the loaded
// libcuda through the runtime (liborbx.so links cudart statically and does not link libcuda).
static int build_tensor_maps(orbx_ctx* c, int policy = 0) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CU(c, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  if (!fn || q != cudaDriverEntryPointSuccess) { c->err = "cuTensorMapEncodeTiled not available"; return ORBX_ERR_CUDA; }
  const Params& P = c->P;
  for (int l = 0; l < P.nlevels; ++l) {
    const LevelInfo& L = P.lv[l];
    // 32-bit elements (a box dimension is limited to 256 elements); columns between w and the pitch are slab padding
    // that no tested pixel's ring reaches, rows beyond h are zero-filled
    cuuint64_t dims[3] = {(cuuint64_t)(L.pitch / 4), (cuuint64_t)L.h, (cuuint64_t)P.maxBatch};
    cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)P.slabBytes};
    cuuint32_t box[3] = {(cuuint32_t)((policy ? L.oBoxW : L.fastBoxW) / 4), (cuuint32_t)(policy ? L.oBoxH : L.fastBoxH), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = ((EncodeFn)fn)(&c->tma.lvl[l], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, P.pyr + L.imgOff, dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { c->err = "cuTensorMapEncodeTiled failed: " + std::to_string((int)r); return ORBX_ERR_CUDA; }
    if (policy == 0) {  // k_describe's patch boxes over the pyramid and the blurred pyramid (independent of the keypoint policy)
      for (int b = 0; b < 2; ++b) {
        cuuint32_t pbox[3] = {(cuuint32_t)kPatchWords, (cuuint32_t)(b ? kPatchRows : kDiscRows), 1};
        r = ((EncodeFn)fn)(b ? &c->patchMaps.blur[l] : &c->patchMaps.pyr[l], CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, (b ? P.blur : P.pyr) + L.imgOff, dims,
                           strides, pbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { c->err = "cuTensorMapEncodeTiled (patch) failed: " + std::to_string((int)r); return ORBX_ERR_CUDA; }
      }
    }
  }
  return ORBX_OK;
}

static int create_impl(orbx_ctx* c, int device, int width, int height, int nfeatures, float scaleFactor,
                       int nlevels, int iniTh, int minTh, int maxBatch) {
  if (width <= 0 || height <= 0 || nfeatures <= 0 || maxBatch <= 0) return ORBX_ERR_ARG;
  int rc = compute_tables(nfeatures, scaleFactor, nlevels, c->T);
  if (rc) return rc;
  Params& P = c->P;
  std::memset(&P, 0, sizeof(P));
  P.W = width; P.H = height; P.nlevels = nlevels; P.nfeatures = nfeatures;
  P.iniTh = iniTh; P.minTh = minTh; P.maxBatch = maxBatch; P.trigMode = 2;
  // the only place the environment is read: schedule options of this context
  auto env_int = [](const char* name, int dflt) {
    const char* v = std::getenv(name);
    if (!v) return dflt;
    return *v ? std::atoi(v) : 1;
  };
  P.optSingleStream = env_int("ORBX_SINGLE_STREAM", 0) != 0;
  P.optRunParts = env_int("ORBX_RUN_PARTS", 0);
  P.optLanes = env_int("ORBX_LANES", 0);
  P.optResizeRows = env_int("ORBX_RESIZE_ROWS", 0);
  P.optMatchSerial = env_int("ORBX_MATCH_SERIAL", 0) != 0;
  P.optMatchPath = env_int("ORBX_MATCH_PATH", 0);
  P.optFrameGraph = env_int("ORBX_FRAME_GRAPH", 1) != 0;
  P.optPyramidChain = env_int("ORBX_PYRAMID_CHAIN", 1) != 0;
  P.optFastTileBudget = std::max(4096, env_int("ORBX_FAST_TILE", 15360));
  std::memcpy(P.umax, c->T.umax, sizeof(P.umax));
  std::vector<StripDesc> strips;
  std::vector<BlurTile> tiles;
  std::vector<u8> rootTab;
  std::vector<int> rootEdges;
  rc = build_geometry(P, c->T, P.optFastTileBudget, strips, tiles, rootTab, rootEdges);
  if (rc) return rc;
  int kp = 0;
  for (int l = 0; l < nlevels; ++l) kp += P.lv[l].quota;
  P.kpStride = std::max(kp, nfeatures);

  int ndev = 0;
  CU(c, cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) { c->err = "no such CUDA device"; return ORBX_ERR_CUDA; }
  c->device = device;
  CU(c, cudaSetDevice(device));
  CU(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->sLane[0] = c->stream;
  for (int i = 0; i < orbx_ctx::kLanes; ++i) {
    if (i) CU(c, cudaStreamCreateWithFlags(&c->sLane[i], cudaStreamNonBlocking));
    CU(c, cudaEventCreateWithFlags(&c->evLane[i], cudaEventDisableTiming));
    CU(c, cudaStreamCreateWithFlags(&c->sSide[i], cudaStreamNonBlocking));
    CU(c, cudaEventCreateWithFlags(&c->evFork[i], cudaEventDisableTiming));
    CU(c, cudaEventCreateWithFlags(&c->evJoin[i], cudaEventDisableTiming));
  }
  CU(c, cudaStreamCreateWithFlags(&c->sH2D, cudaStreamNonBlocking));
  CU(c, cudaStreamCreateWithFlags(&c->sD2H, cudaStreamNonBlocking));
  CU(c, cudaHostAlloc((void**)&c->hStatus, 2 * (size_t)maxBatch * sizeof(int), cudaHostAllocDefault));
  for (int i = 0; i < 2; ++i) CU(c, cudaEventCreateWithFlags(&c->evDone[i], cudaEventDisableTiming));
  CU(c, cudaEventCreate(&c->evStart));
  CU(c, cudaEventCreate(&c->evStop));
  for (int i = 0; i < 10; ++i) CU(c, cudaEventCreate(&c->evStage[i]));

  const size_t B = (size_t)maxBatch;
  StripDesc* dStrips; BlurTile* dTiles; u8* dRootTab; int* dRootEdges;
  if ((rc = dupload(c, &dStrips, strips))) return rc;
  if ((rc = dupload(c, &dTiles, tiles))) return rc;
  if ((rc = dupload(c, &dRootTab, rootTab))) return rc;
  if ((rc = dupload(c, &dRootEdges, rootEdges))) return rc;
  P.strips = dStrips; P.blurTiles = dTiles; P.rootTab = dRootTab; P.rootEdges = dRootEdges;
  c->dResizeTabs.assign(nlevels, nullptr);
  c->hTabPtrs = new const int*[nlevels];
  c->hTabPtrs[0] = nullptr;
  for (int l = 1; l < nlevels; ++l) {
    std::vector<int> tab;
    build_resize_tab(P.lv[l - 1].w, P.lv[l - 1].h, P.lv[l].w, P.lv[l].h, tab);
    if ((rc = dupload(c, &c->dResizeTabs[l], tab))) return rc;
    c->hTabPtrs[l] = c->dResizeTabs[l];
  }
  if ((rc = dalloc(c, &P.pyr, B * P.slabBytes + 256))) return rc;   // + slack: aligned word reads may run a few bytes past a row
  if ((rc = dalloc(c, &P.blur, B * P.slabBytes + 256))) return rc;
  if ((rc = dalloc(c, &P.candKey, B * P.candTotal))) return rc;
  if ((rc = dalloc(c, &P.candXY, B * P.candTotal))) return rc;
  if ((rc = dalloc(c, &P.candNode, B * P.candTotal))) return rc;
  if ((rc = dalloc(c, &P.candCount, B * kMaxLevels))) return rc;
  if ((rc = dalloc(c, &P.nodeBox, B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeCount, B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeSlot, B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeBest, B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeAux, 4 * B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeList, B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeScan, B * P.nodeTotal))) return rc;
  if ((rc = dalloc(c, &P.nodeSort, B * P.nodeTotal))) return rc;
  const size_t K = B * (size_t)P.kpStride;
  if ((rc = dalloc(c, &P.lvlXY, K))) return rc;
  if ((rc = dalloc(c, &P.lvlResp, K))) return rc;
  if ((rc = dalloc(c, &P.lvlAngle, K))) return rc;
  if ((rc = dalloc(c, &P.lvlDesc, K * 32))) return rc;
  if ((rc = dalloc(c, &P.lvlCount, B * kMaxLevels))) return rc;
  if ((rc = dalloc(c, &P.status, B))) return rc;
  if ((rc = dalloc(c, &P.gridBar, 1))) return rc;
  if (cudaDeviceGetAttribute(&P.smCount, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return ORBX_ERR_CUDA;
  if ((rc = dalloc(c, &P.outKp, K))) return rc;
  if ((rc = dalloc(c, &P.outDesc, K * 32))) return rc;
  if ((rc = dalloc(c, &P.outN, B))) return rc;
  if ((rc = dalloc(c, &P.outMono, B))) return rc;
  if ((rc = dalloc(c, &P.keysUn, K))) return rc;
  if ((rc = dalloc(c, &P.gridOff, B * (kGridCells + 1)))) return rc;
  if ((rc = dalloc(c, &P.gridIdx, K))) return rc;
  const size_t nPairs = (B + 1) / 2;
  if ((rc = dalloc(c, &c->dMatches12, nPairs * P.kpStride))) return rc;
  if ((rc = dalloc(c, &c->dNMatches, nPairs))) return rc;
  if ((rc = dalloc(c, &c->dCounters, nPairs * 3))) return rc;
  if ((rc = dalloc(c, &c->dMatchScratch, nPairs * 3 * P.kpStride))) return rc;
  auto bf_alloc = [&](BfScratch& b, size_t np, size_t stride) {
    int r;
    if ((r = dalloc(c, &b.cand, np * stride))) return r;
    if ((r = dalloc(c, &b.candXY, np * stride))) return r;
    if ((r = dalloc(c, &b.qIdx, np * stride))) return r;
    if ((r = dalloc(c, &b.counts, np * 4))) return r;
    return dalloc(c, &b.topk, np * stride * 8);
  };
  if ((rc = bf_alloc(c->dBf, nPairs, (size_t)P.kpStride))) return rc;
  if ((rc = bf_alloc(c->hfBf, 1, (size_t)P.kpStride))) return rc;
  // host-frame staging (2 frames)
  c->hfCap = P.kpStride;
  {
    const size_t cap = (size_t)c->hfCap;
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t oKeys = 0, oDesc = oKeys + up(2 * cap * sizeof(orbx_keypoint)), oOff = oDesc + up(2 * cap * 32),
                 oIdx = oOff + up(2 * (size_t)(kGridCells + 1) * 4), oN = oIdx + up(2 * cap * 4), inBytes = oN + 256;
    const size_t oNM = inBytes, oCnt = oNM + 4, oM = oNM + 16, outBytes = 16 + cap * 4;
    if ((rc = dalloc(c, &c->hfBlock, inBytes + outBytes))) return rc;
    c->hfKeys = reinterpret_cast<orbx_keypoint*>(c->hfBlock + oKeys);
    c->hfDesc = c->hfBlock + oDesc;
    c->hfGridOff = reinterpret_cast<u32*>(c->hfBlock + oOff);
    c->hfGridIdx = reinterpret_cast<u32*>(c->hfBlock + oIdx);
    c->hfN = reinterpret_cast<int*>(c->hfBlock + oN);
    c->hfNMatches = reinterpret_cast<int*>(c->hfBlock + oNM);
    c->hfCounters = reinterpret_cast<int*>(c->hfBlock + oCnt);
    c->hfMatches = reinterpret_cast<int*>(c->hfBlock + oM);
    c->hfInBytes = inBytes; c->hfOutBytes = outBytes;
    CU(c, cudaHostAlloc((void**)&c->hHfIn, inBytes, cudaHostAllocDefault));
    CU(c, cudaHostAlloc((void**)&c->hHfOut, outBytes, cudaHostAllocDefault));
    std::memset(c->hHfIn, 0, inBytes);
  }
  if ((rc = dalloc(c, &c->hfScratch, 3 * (size_t)c->hfCap))) return rc;
  if ((rc = dalloc(c, &c->dIngest, B * (size_t)width * height))) return rc;
  if ((rc = build_tensor_maps(c))) return rc;
  if ((rc = configure_fast(P))) { c->err = "k_fast: shared memory / occupancy configuration failed"; return rc; }
  if ((rc = configure_quadtree())) { c->err = "k_quadtree: shared memory configuration failed"; return rc; }
  c->fastSet[0].strips = P.strips; c->fastSet[0].nStrips = P.nStrips; c->fastSet[0].tileMax = P.fastTileMax;
  c->fastSet[0].segW = P.fastSegW; c->fastSet[0].segP = P.fastSegP; c->fastSet[0].grid = P.fastGrid; c->fastSet[0].tma = c->tma;
  c->fastSet[0].ready = 1; c->fastSet[1].ready = 0;
  CU(c, cudaMemset(P.status, 0, B * sizeof(int)));
  CU(c, cudaMemset(P.outN, 0, B * sizeof(int)));
  CU(c, cudaMemset(P.pyr, 0, B * P.slabBytes));
  CU(c, cudaMemset(P.blur, 0, B * P.slabBytes));
  CU(c, cudaDeviceSynchronize());
  return ORBX_OK;
}

extern "C" orbx_ctx* orbx_create(int device, int width, int height, int nfeatures, float scale_factor, int nlevels,
                                 int ini_th_fast, int min_th_fast, int max_batch, int* status) {
  orbx_ctx* c = new (std::nothrow) orbx_ctx();
  if (!c) { if (status) *status = ORBX_ERR_ARG; return nullptr; }
  c->stream = nullptr; c->sH2D = c->sD2H = nullptr;
  for (int i = 0; i < orbx_ctx::kLanes; ++i) { c->sLane[i] = c->sSide[i] = nullptr; c->evFork[i] = c->evJoin[i] = c->evLane[i] = nullptr; }
  c->dIngest = nullptr; c->dIngestColor = nullptr; c->hStatus = nullptr; c->hTabPtrs = nullptr; c->launches = 0; c->cameraSet = 0; c->lastBatch = 0; c->dScore = nullptr; c->scoreBytes = 0;
  for (int i = 0; i < 2; ++i) { c->flight[i].active = 0; c->flight[i].n = 0; c->flight[i].chunk = 0; c->flight[i].status = 0; c->evDone[i] = nullptr; }
  c->nextTicket = 0;
  for (int i = 0; i < 2; ++i) { c->fg[i].exec = nullptr; c->fg[i].epoch = -1; }
  c->graphEpoch = 0; c->hFrameIn = nullptr; c->hFrameOut = nullptr; c->hHfIn = nullptr; c->hHfOut = nullptr; c->hfBlock = nullptr;
  c->stageTiming = 0; c->evStart = c->evStop = nullptr;
  for (int i = 0; i < 10; ++i) c->evStage[i] = nullptr;
  std::memset(c->stageMs, 0, sizeof(c->stageMs));
  int rc = create_impl(c, device, width, height, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast, max_batch);
  if (status) *status = rc;
  if (rc != ORBX_OK) {
    if (rc == ORBX_ERR_CUDA) std::fprintf(stderr, "orbx_create: %s\n", c->err.c_str());
    orbx_destroy(c);
    return nullptr;
  }
  return c;
}

extern "C" void orbx_destroy(orbx_ctx* c) {
  if (!c) return;
  if (c->stream) cudaStreamSynchronize(c->stream);
  for (void* p : c->allocs) cudaFree(p);
  if (c->dScore) cudaFree(c->dScore);
  if (c->evStart) cudaEventDestroy(c->evStart);
  if (c->evStop) cudaEventDestroy(c->evStop);
  for (int i = 0; i < 10; ++i) if (c->evStage[i]) cudaEventDestroy(c->evStage[i]);
  for (cudaEvent_t e : c->evChunk) cudaEventDestroy(e);
  if (c->hStatus) cudaFreeHost(c->hStatus);
  for (int i = 0; i < 2; ++i) if (c->fg[i].exec) cudaGraphExecDestroy(c->fg[i].exec);
  if (c->hFrameIn) cudaFreeHost(c->hFrameIn);
  if (c->hFrameOut) cudaFreeHost(c->hFrameOut);
  if (c->hHfIn) cudaFreeHost(c->hHfIn);
  if (c->hHfOut) cudaFreeHost(c->hHfOut);
  for (int i = 0; i < 2; ++i) if (c->evDone[i]) cudaEventDestroy(c->evDone[i]);
  for (int i = 0; i < orbx_ctx::kLanes; ++i) {
    if (c->sSide[i]) cudaStreamDestroy(c->sSide[i]);
    if (c->evFork[i]) cudaEventDestroy(c->evFork[i]);
    if (c->evJoin[i]) cudaEventDestroy(c->evJoin[i]);
    if (c->evLane[i]) cudaEventDestroy(c->evLane[i]);
    if (i && c->sLane[i]) cudaStreamDestroy(c->sLane[i]);
  }
  if (c->sH2D) cudaStreamDestroy(c->sH2D);
  if (c->sD2H) cudaStreamDestroy(c->sD2H);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete[] c->hTabPtrs;
  delete c;
}

extern "C" const char* orbx_last_error(const orbx_ctx* c) { return c ? c->err.c_str() : "null context"; }
extern "C" int orbx_set_trig_mode(orbx_ctx* c, int mode) {
  if (!c || mode < 0 || mode > 2) return ORBX_ERR_ARG;
  c->P.trigMode = mode;
  ++c->graphEpoch;
  return ORBX_OK;
}
// ComputeKeyPointsOctTree (0, what the reference's operator() calls) or ComputeKeyPointsOld (1, Features/ORBextractor.cpp:1207-1496)
extern "C" int orbx_set_keypoint_policy(orbx_ctx* c, int policy) {
  if (!c || (policy != 0 && policy != 1)) return ORBX_ERR_ARG;
  CU(c, cudaSetDevice(c->device));
  int rc = drain_flights(c);
  if (rc) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  Params& P = c->P;
  orbx_ctx::FastSet& S = c->fastSet[policy];
  if (!S.ready) {  // first use of ComputeKeyPointsOld: its cell grid, strips, tensor maps and scratch
    std::vector<StripDesc> strips;
    int cellTotal = 0;
    if ((rc = build_strips_old(P, c->T, P.optFastTileBudget, strips, S.tileMax, S.segW, S.segP, cellTotal))) return rc;
    if (strips.empty()) return ORBX_ERR_ARG;
    StripDesc* d = nullptr;
    if ((rc = dupload(c, &d, strips))) return rc;
    S.strips = d; S.nStrips = (int)strips.size();
    if ((rc = dalloc(c, &P.oldSorted, (size_t)P.maxBatch * P.candTotal))) return rc;
    P.oldCellTotal = cellTotal;
    const TmaMaps keep = c->tma;
    if ((rc = build_tensor_maps(c, 1))) { c->tma = keep; return rc; }
    S.tma = c->tma;
    c->tma = keep;
    Params Q = P;
    Q.strips = S.strips; Q.nStrips = S.nStrips; Q.fastTileMax = S.tileMax; Q.fastSegW = S.segW; Q.fastSegP = S.segP;
    if ((rc = configure_fast(Q))) { c->err = "k_fast: shared memory / occupancy configuration failed"; return rc; }
    S.grid = Q.fastGrid;
    S.ready = 1;
  }
  P.strips = S.strips; P.nStrips = S.nStrips; P.fastTileMax = S.tileMax; P.fastSegW = S.segW; P.fastSegP = S.segP; P.fastGrid = S.grid;
  c->tma = S.tma;
  P.kpPolicy = policy;
  ++c->graphEpoch;
  return ORBX_OK;
}

extern "C" int orbx_set_option(orbx_ctx* c, const char* name, int value) {
  if (!c || !name) return ORBX_ERR_ARG;
  const std::string n(name);
  if (n == "single_stream") c->P.optSingleStream = value != 0;
  else if (n == "run_parts") c->P.optRunParts = value;
  else if (n == "lanes") c->P.optLanes = value;
  else if (n == "resize_rows") c->P.optResizeRows = value;
  else if (n == "match_serial") c->P.optMatchSerial = value != 0;
  else if (n == "match_path") c->P.optMatchPath = value;
  else if (n == "frame_graph") c->P.optFrameGraph = value != 0;
  else if (n == "qt_tight") c->P.optQtTight = value != 0;
  else if (n == "pyramid_chain") c->P.optPyramidChain = value != 0;
  else return ORBX_ERR_ARG;
  ++c->graphEpoch;
  return ORBX_OK;
}
extern "C" void* orbx_get_stream(orbx_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int orbx_last_launch_count(const orbx_ctx* c) { return c ? c->launches : 0; }

static int copy_tables(const HostTables& T, int nlevels, float* scale, float* inv, float* s2, float* is2, int* npl,
                       int* umax16) {
  for (int i = 0; i < nlevels; ++i) {
    if (scale) scale[i] = T.scale[i];
    if (inv) inv[i] = T.invScale[i];
    if (s2) s2[i] = T.sigma2[i];
    if (is2) is2[i] = T.invSigma2[i];
    if (npl) npl[i] = T.perLevel[i];
  }
  if (umax16) std::memcpy(umax16, T.umax, 16 * sizeof(int));
  return nlevels;
}
extern "C" int orbx_get_level_tables(const orbx_ctx* c, float* scale, float* inv_scale, float* sigma2,
                                     float* inv_sigma2, int* n_per_level, int* umax16) {
  if (!c) return ORBX_ERR_ARG;
  return copy_tables(c->T, c->P.nlevels, scale, inv_scale, sigma2, inv_sigma2, n_per_level, umax16);
}
extern "C" int orbx_compute_level_tables(int nfeatures, float scale_factor, int nlevels, float* scale,
                                         float* inv_scale, float* sigma2, float* inv_sigma2, int* n_per_level,
                                         int* umax16) {
  HostTables T;
  int rc = compute_tables(nfeatures, scale_factor, nlevels, T);
  if (rc) return rc;
  return copy_tables(T, nlevels, scale, inv_scale, sigma2, inv_sigma2, n_per_level, umax16);
}
extern "C" int orbx_get_level_size(const orbx_ctx* c, int level, int* w, int* h) {
  if (!c || level < 0 || level >= c->P.nlevels) return ORBX_ERR_ARG;
  if (w) *w = c->P.lv[level].w;
  if (h) *h = c->P.lv[level].h;
  return ORBX_OK;
}

// DBoW2::FORB::distance (Thirdparty/DBoW2/src/FORB.cpp:77-100): 4 x 64-bit XOR + popcount.
extern "C" int orbx_descriptor_distance(const uint8_t a[32], const uint8_t b[32]) {
  uint64_t x[4], y[4];
  std::memcpy(x, a, 32);
  std::memcpy(y, b, 32);
  int d = 0;
  for (int i = 0; i < 4; ++i) d += __builtin_popcountll(x[i] ^ y[i]);
  return d;
}
