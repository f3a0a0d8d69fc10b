// Shared definitions of the orbx kernels (sm_100a only).
//
// HBM layout of a context (one extractor configuration, `maxBatch` frame slots):
//   pyr / blur   u8   [slot][slabBytes]       level l of a slot at +lv[l].imgOff, rows of lv[l].pitch bytes
//   candKey      u64  [slot][candTotal]       (response << 32) | ~order   (order = reference emission order)
//   candXY       u32  [slot][candTotal]       x | y << 16, relative to (16, 16) as in the reference
//   candNode     u32  [slot][candTotal]       current quadtree node of the candidate
//   candCount    u32  [slot][nlevels]
//   node tables  per (slot, level), lv[l].nodeCap entries
//   lvlKp*       per (slot, level) quota-sized slots: xy, response, angle, descriptor
//   out*         per slot: packed cv::KeyPoint records + descriptors in the reference's order
//   frame store  per slot: undistorted keys, 64x48 CSR grid
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbx.h"
#include "stdsort_emul.cuh"

namespace orbx {

typedef unsigned char u8;
typedef unsigned int u32;
typedef unsigned long long u64;

constexpr int kMaxLevels = ORBX_MAX_LEVELS;
constexpr int kEdge = 19;       // EDGE_THRESHOLD, Features/ORBextractor.cpp:76
constexpr int kHalfPatch = 15;  // HALF_PATCH_SIZE, :74
constexpr int kPatch = 31;      // PATCH_SIZE, :72
constexpr int kMinBorder = 16;  // EDGE_THRESHOLD - 3, :1056
constexpr int kGridCols = ORBX_GRID_COLS, kGridRows = ORBX_GRID_ROWS, kGridCells = ORBX_GRID_CELLS;
#ifndef ORBX_BLUR_TH
#define ORBX_BLUR_TH 124
#endif
constexpr int kBlurTW = 120, kBlurTH = ORBX_BLUR_TH;  // k_blur: output columns x rows per warp strip
// cell index of an interior column by multiply and shift: exact for xb * wCell < 2^22; xb * cellMul stays below 2^32 (xb <= 244 when
// a strip holds several cells, xb < wCell <= 480 otherwise)
constexpr int kCellShift = 22;
#ifndef ORBX_FAST_WARPS
#define ORBX_FAST_WARPS 8
#endif
constexpr int kFastWarps = ORBX_FAST_WARPS;   // warps of a k_fast CTA (its per-warp list segments are sized on the host)
constexpr int kMaxCellDim = 80;  // cell ROI (wCell + 6) is < 76 for every image size

struct LevelInfo {
  int w, h, pitch;    // unpadded size, row pitch in bytes (multiple of 128)
  u32 imgOff;         // byte offset of the level in a slot's slab
  int nCols, nRows, wCell, hCell;  // FAST cell grid, :1066-1075
  int fastBoxW, fastBoxH;  // TMA box of k_fast at this level: bytes per staged row (an odd multiple of 16), rows
  // ComputeKeyPointsOld's own cell grid (Features/ORBextractor.cpp:1222-1251), filled when that policy is selected
  int oCols, oRows, oCellW, oCellH, oNfCell;  // levelCols, levelRows, cellW, cellH, nfeaturesCell
  int oBoxW, oBoxH;        // TMA box of k_fast for that grid
  int quota;          // mnFeaturesPerLevel[l]
  int candCap;        // capacity of the candidate arrays of this level
  u32 candOff;        // first candidate of this level in a slot's candidate arrays
  int nodeCap;        // capacity of the quadtree node tables
  u32 nodeOff;
  int nIni;           // number of root nodes, :706
  u32 rootTabOff;     // x -> root index table (u8), precomputed with the reference's float expression
  u32 rootEdgeOff;    // nIni + 1 root box edges (int)(hX * i), :719-724
  int kpOff;          // first keypoint slot of this level (prefix sum of quotas)
  float scale;        // mvScaleFactor[l]
  float patchSize;    // (float)(int)(31 * mvScaleFactor[l]), :1165
  int distW, distH;   // maxBorderX - minBorderX, maxBorderY - minBorderY
};

struct StripDesc {    // one work item of k_fast: a run of consecutive FAST cells of one cell row, Features/ORBextractor.cpp:1078-1141.
                      // Everything a CTA needs about the strip and its level, derived on the host (48 bytes, three 16-byte loads).
  short level, nCells;
  short tmaX, y0;         // tile origin: x in 32-bit elements (a 16-byte aligned column <= x0 - 1), first row of the cells' ROIs
  short SB;               // staged row stride in bytes = lv[level].fastBoxW
  short bLo, bHi;         // interior byte columns [bLo, bHi) of a staged row: level columns [x0 + 3, x1 - 3) of the cells' ROIs [x0, x1);
                          // cell j of the strip tests the columns [bLo + j * wCell, min(bLo + (j + 1) * wCell, bHi))
  short nR;               // interior rows: tile rows [3, 3 + nR)
  short wCell;
  short sx, sy;           // j0 * wCell, i * hCell: shift applied to ROI-relative coordinates (:1134-1135)
  short nIniM1;           // lv[level].nIni - 1
  int tileBytes;          // box bytes of the level: what the tile copy delivers; bits 24..27: retryMax (a cell with at most that many
                          // keypoints at iniThFAST runs again at minThFAST: 0 in ComputeKeyPointsOctTree, 3 in ComputeKeyPointsOld);
                          // bits 28..31: bits per coordinate of the emission-order key (7 or 9)
  u32 cellMul;            // (xb * cellMul) >> kCellShift == xb / wCell for every interior column xb of a strip
  int index0;             // i * nCols + j0: row-major index of the strip's first cell
  u32 candOff;            // lv[level].candOff, candCap, rootTabOff
  int candCap;
  u32 rootTabOff;
};
static_assert(sizeof(StripDesc) == 48, "StripDesc is loaded as three 16-byte words");

struct BlurTile {
  short level, pad;
  short x0, y0;
};

// TMA descriptors (cuTensorMapEncodeTiled, built once in orbx_create): level l of the pyramid slab as a
// u32 tensor {x: pitch / 4, y: h, slot: maxBatch} with strides {pitch, slabBytes}; box = lv[l].fastBoxW / 4 x lv[l].fastBoxH x 1
// (32-bit elements: a box dimension is limited to 256 elements, and a strip row is wider than 256 bytes).
struct TmaMaps {
  CUtensorMap lvl[kMaxLevels];
};
// k_describe's keypoint patches: the same tensors over P.pyr (IC_Angle, 31 rows) and P.blur (steered BRIEF, 37 rows) with
// boxes of kPatchWords words per row.  A box must start on a 16-byte boundary in global memory (anything else raises Warp
// Illegal Instruction), so a patch starts at the 16-byte column at or below x - 18 and is 64 bytes wide.
constexpr int kPatchRows = 37, kDiscRows = 31, kPatchWords = 16;
struct PatchMaps {
  CUtensorMap pyr[kMaxLevels], blur[kMaxLevels];
};
__device__ __forceinline__ u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }

// Inclusive prefix sum of one value per thread over a CTA of NT threads (a multiple of 32, at most 1024): shuffles inside the
// warps, the warp totals through part[NT / 32]; two CTA barriers.  *total receives the sum over the CTA.
template <int NT>
__device__ __forceinline__ int cta_prefix_sum(int v, int* part, int* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) part[w] = v;
  __syncthreads();
  int base = 0, sum = 0;
#pragma unroll
  for (int i = 0; i < NT / 32; ++i) {
    const int p = part[i];
    if (i < w) base += p;
    sum += p;
  }
  __syncthreads();  // part[] may be reused
  *total = sum;
  return base + v;
}

struct Params {
  int W, H, nlevels, nfeatures, iniTh, minTh, maxBatch;
  int kpStride;        // keypoint slots per frame = max(nfeatures, sum of quotas)
  int trigMode;
  int nStrips, nBlurTiles;
  int slot0;               // first slot of this launch inside the context's slot arrays (TMA z origin; pointers below are pre-shifted)
  int fastTileMax;         // k_fast shared memory: bytes of one staged tile (largest level box, multiple of 128)
  int fastSegW, fastSegP;  // per-warp capacities (entries) of the word and the pixel segments: the worst case of any strip
  int fastGrid;            // persistent CTAs of a full k_fast launch (SMs x resident CTAs per SM)
  u32 slabBytes, candTotal, nodeTotal;
  int umax[16];
  LevelInfo lv[kMaxLevels];
  // device pointers
  u8* pyr;
  u8* blur;
  u64* candKey;
  u32* candXY;
  u32* candNode;
  u32* candCount;      // [slot][kMaxLevels]
  const StripDesc* strips;
  int kpPolicy;            // 0: ComputeKeyPointsOctTree (default), 1: ComputeKeyPointsOld (orbx_set_keypoint_policy)
  u64* oldSorted;          // [slot][candTotal]  ComputeKeyPointsOld: candidates in the reference's order (response << 48 | y << 16 | x ...)
  int* oldCell;            // [slot][oldCellTotal * 4]  per cell: count, start, nToRetain, flags
  int oldCellTotal;        // cells of all levels (per slot)
  const BlurTile* blurTiles;
  const u8* rootTab;
  // quadtree node tables [slot][nodeTotal]
  short4* nodeBox;     // ulx, uly, urx, bry
  int* nodeCount;
  int* nodeSlot;       // split slot of the node in the current pass, -1 if not being split
  u64* nodeBest;
  int* nodeAux;        // 4 ints per split slot: child counts, then child ids, of the current pass
  int* nodeList;       // split list of the current pass (node ids in visiting order)
  int* nodeScan;       // scan scratch
  SortElem* nodeSort;  // (count, UL.x) pairs for the std::sort replay
  const int* rootEdges;
  // per-level keypoints [slot][nfeatures]
  u32* lvlXY;          // x | y << 16, level coordinates
  float* lvlResp;
  float* lvlAngle;
  u8* lvlDesc;         // 32 bytes each
  int* lvlCount;       // [slot][kMaxLevels]
  int* status;         // [slot]  sticky error flags of the slot (ORBX_ERR_*), 0 = ok
  u32* gridBar;        // arrival counter of k_pyramid_chain's grid barrier
  int smCount;         // SMs of the context's device
  // packed outputs [slot][nfeatures]
  orbx_keypoint* outKp;
  u8* outDesc;
  int* outN;           // [slot]
  int* outMono;        // [slot]
  // frame store [slot]
  orbx_keypoint* keysUn;
  u32* gridOff;        // [slot][kGridCells + 1]
  u32* gridIdx;        // [slot][nfeatures]
  // camera
  float K4[4], D4[4];
  int bounds[4];
  float gridWInv, gridHInv;
  int undistort;       // mDistCoef(0) != 0
  // host-side schedule options (kernels ignore them): read ONCE in orbx_create from the environment (ORBX_SINGLE_STREAM,
  // ORBX_RUN_PARTS, ORBX_LANES, ORBX_RESIZE_ROWS, ORBX_MATCH_SERIAL, ORBX_FAST_TILE) or set with orbx_set_option
  int optSingleStream;   // profiling aid: whole-batch launches in order on one stream
  int optRunParts;       // pieces of orbx_stream_run (0 = default)
  int optLanes;          // compute lanes of the chunked host path (0 = default)
  int optResizeRows;     // rows per warp of k_resize (0 = default)
  int optMatchSerial;    // test hook: force the one-warp-per-pair matcher
  int optMatchPath;      // 0 = choose by window and capacity, 1 = k_match_cta, 2 = k_match, 3 = brute force (match_bf.cu)
  int optFastTileBudget; // bytes of one staged k_fast tile (creation time only)
  int optFrameGraph;     // single-frame calls run through a CUDA graph (default 1)
  int optPyramidChain;   // batches of at most 2 frames build the pyramid in one launch (default 1)
  int optQtTight;        // test hook: k_quadtree squeezes its node table at the smallest size its bound allows (results unchanged)
};

__device__ __forceinline__ const u8* level_ptr(const u8* base, const Params& P, int slot, int l) {
  return base + (size_t)slot * P.slabBytes + P.lv[l].imgOff;
}
__device__ __forceinline__ u8* level_ptr(u8* base, const Params& P, int slot, int l) {
  return base + (size_t)slot * P.slabBytes + P.lv[l].imgOff;
}
__device__ __forceinline__ int reflect101(int i, int len) {
  // BORDER_REFLECT_101, single reflection is enough for |overhang| < len
  if (i < 0) i = -i;
  if (i >= len) i = 2 * (len - 1) - i;
  return i;
}

// launchers (one per .cu file); all enqueue on `st` and return the number of kernel launches
int launch_ingest(const Params& P, const u8* dense, int nSlots, cudaStream_t st);
int launch_ingest_color(const Params& P, const u8* dense, int rgb, int nSlots, cudaStream_t st);
int launch_pyramid(const Params& P, const int* const* dTabs, int nSlots, cudaStream_t st);
int launch_blur(const Params& P, int nSlots, cudaStream_t st);
int launch_fast(const Params& P, const TmaMaps& M, int nSlots, cudaStream_t st);
int launch_quadtree(const Params& P, int nSlots, cudaStream_t st);
int configure_quadtree();
int launch_retain_old(const Params& P, int nSlots, cudaStream_t st);
int launch_describe(const Params& P, const PatchMaps& M, int nSlots, int lap0, int lap1, cudaStream_t st);
void launch_sincosf_probe(const float* y, long long n, float* sn, float* cs, cudaStream_t st);
int launch_glue(const Params& P, int nSlots, cudaStream_t st);
// scratch of the brute-force matcher (match_bf.cu), per pair; topk == nullptr: not available
struct BfScratch {
  u32* cand;        // [pair][stride]  level-0 keys of F2 in visiting order: index | cell ix << 16 | cell iy << 22
  float2* candXY;   // [pair][stride]  their undistorted positions
  int* qIdx;        // [pair][stride]  level-0 keys of F1, ascending
  int* counts;      // [pair][4]       number of candidates, number of queries
  u32* topk;        // [pair][stride][8]  per query: 7 keys (distance << 20 | visiting position), info word
};
struct MatchArgs {
  const orbx_keypoint* keysUn;  // [slot][nfeatures]
  const u8* desc;
  const u32* gridOff;
  const u32* gridIdx;
  const int* n;                 // [slot]
  int stride;                   // nfeatures
  int* matches12;               // [pair][stride]
  int* nmatches;                // [pair]
  int* counters;                // [pair][3]
  int* scratch;                 // [pair][2*stride + ...]
  BfScratch bf;                 // already advanced to the first pair of the launch
};
int launch_match(const Params& P, const MatchArgs& A, int nPairs, int pairStep, float nnratio, int checkOri,
                 int window, cudaStream_t st);
int launch_match_bf(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, float nnratio, int checkOri,
                    int window, cudaStream_t st);
int launch_bf_prep(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, cudaStream_t st);
int launch_bf_rows(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, int window, int q0, int q1,
                   cudaStream_t st);
int launch_bf_resolve(const Params& P, const MatchArgs& A, const BfScratch& B, int nPairs, int pairStep, float nnratio, int checkOri,
                      int window, cudaStream_t st);
int launch_hamming_topk(const u8* a, int nA, const u8* b, int nB, u32* topk, cudaStream_t st);

// Initializer::CheckHomography / CheckFundamental over hypothesis sets (score.cu)
struct ScoreArgs {
  const orbx_keypoint* keysUn;  // [slot][stride]
  const int* n;                 // [slot]
  int stride, pairStep;         // frames of pair p live in slots p * pairStep and p * pairStep + 1
  const int* matches12;         // [pair][stride]
  const float* hypA;            // [pair][nHyp][9]  H21 or F21, row-major
  const float* hypB;            // [pair][nHyp][9]  H12 (homography only)
  int nHyp;
  float invSigmaSquare;
  float* scores;                // [pair][nHyp]
  int* best;                    // [pair]  -1: no hypothesis scored above 0
  int* nList;                   // [pair]  N = length of the match list
  u8* inlAll;                   // nullable, [pair][nHyp][N] (packed with the pair's own N; single-pair calls)
  u8* inlBest;                  // nullable, [pair][stride]: inlier flags of the selected hypothesis, 0 beyond N
};
int launch_score_models(const ScoreArgs& S, int nPairs, bool fundamental, cudaStream_t st);

}  // namespace orbx
