// Host-side context of liborbx.so: owns the stream, every device allocation and the host tables.
#pragma once
#include <string>
#include <vector>

#include "common.cuh"

namespace orbx {
void launch_hamming(const u8* a, const u8* b, int n, int* out, cudaStream_t st);

// ORBextractor::ORBextractor host arithmetic (Features/ORBextractor.cpp:492-595)
struct HostTables {
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> perLevel;
  int umax[16];
};
int compute_tables(int nfeatures, float scaleFactor, int nlevels, HostTables& t);
}  // namespace orbx

#define CU(ctx, call)                                                                       \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                      \
      return ORBX_ERR_CUDA;                                                                 \
    }                                                                                       \
  } while (0)

struct orbx_ctx {
  orbx::Params P;
  orbx::HostTables T;
  orbx::PatchMaps patchMaps;       // per-level tensor maps over P.pyr and P.blur with k_describe's patch boxes
  orbx::TmaMaps tma;               // per-level tensor maps over P.pyr (k_fast tile staging) of the ACTIVE keypoint policy
  struct FastSet {                 // what k_fast needs per keypoint policy ([0] ComputeKeyPointsOctTree, [1] ComputeKeyPointsOld)
    const orbx::StripDesc* strips;
    int nStrips, tileMax, segW, segP, grid, ready;
    orbx::TmaMaps tma;
  } fastSet[2];
  int device;
  cudaStream_t stream;
  static constexpr int kLanes = 4;
  cudaStream_t sLane[kLanes];      // compute lanes: [0] = stream; chunks of orbx_stream_process / parts of orbx_stream_run rotate over them
  cudaStream_t sSide[kLanes];      // blur runs here, beside the quadtree of the same lane
  cudaEvent_t evFork[kLanes], evJoin[kLanes], evLane[kLanes];
  cudaStream_t sH2D, sD2H;         // copy streams of the pipelined host path (orbx_stream_process)
  std::vector<cudaEvent_t> evChunk; // 5 per chunk: images landed, kernels done, staging consumed, results on the host, status on the host
  orbx::u8* dIngest;               // dense [maxBatch][H][W] staging of uploaded frames
  orbx::u8* dIngestColor;          // dense [maxBatch][H][W][3], allocated on the first colour upload
  int* hStatus;                    // pinned [2][maxBatch]: one block per step in flight
  struct Flight { int active, n, chunk, status; } flight[2];  // steps of orbx_stream_process_async not yet waited for; status: the
                                                              // step's worst device-side status once it has been drained (0 = ok / not known yet)
  cudaEvent_t evDone[2];
  int nextTicket;
  std::vector<void*> allocs;       // every cudaMalloc, freed in orbx_destroy
  std::vector<int*> dResizeTabs;   // per level (entry 0 unused)
  const int** hTabPtrs;            // host array of device pointers for launch_pyramid
  std::string err;
  int launches;
  int cameraSet;
  int lastBatch;
  // matcher outputs for the stream path [pair]
  int* dMatches12;
  int* dNMatches;
  int* dCounters;
  int* dMatchScratch;
  orbx::BfScratch dBf;      // brute-force matcher scratch of the stream path ([pair] arrays)
  orbx::BfScratch hfBf;     // the same for host-frame calls (one pair)
  // staging store for host-frame calls (2 frames): keysUn, desc, gridOff, gridIdx, n
  orbx_keypoint* hfKeys;
  orbx::u8* hfDesc;
  orbx::u32* hfGridOff;
  orbx::u32* hfGridIdx;
  int* hfN;
  int* hfMatches;   // matches12 + nmatches + counters + scratch
  int* hfNMatches;
  int* hfCounters;
  int* hfScratch;
  int hfCap;
  // the staging store is ONE device block, inputs first (keys | desc | gridOff | gridIdx | n), then the results (nmatches | counters |
  // matches12), with a pinned host mirror of each part: a host-frame call is one H2D copy, its kernels, one D2H copy
  orbx::u8* hfBlock;
  orbx::u8* hHfIn;    // pinned, hfInBytes
  orbx::u8* hHfOut;   // pinned, hfOutBytes
  size_t hfInBytes, hfOutBytes;
  // Initializer scoring scratch (grown on demand): hypotheses, scores, best, list lengths, inlier flags
  orbx::u8* dScore;
  size_t scoreBytes;
  // single-frame latency path: the whole launch sequence of one frame (H2D of the image from the pinned staging buffer,
  // extraction, optionally the Frame glue, D2H of every result into the pinned result block) as ONE CUDA graph, built on
  // first use and rebuilt when what it bakes in changes (lapping area, camera, trig mode, schedule options)
  struct FrameGraph {
    cudaGraphExec_t exec;
    int lap0, lap1, withGlue, epoch;
  } fg[2];                       // [0] extraction only (orbx_extract), [1] extraction + glue (orbx_frame_construct)
  int graphEpoch;                // bumped by every setter a graph depends on
  orbx::u8* hFrameIn;            // pinned [H][W]
  orbx::u8* hFrameOut;           // pinned result block, see frame_out_layout()
  // timing
  cudaEvent_t evStart, evStop;
  int stageTiming;
  cudaEvent_t evStage[10];
  float stageMs[8];
};

int drain_flights(orbx_ctx* c);
