// ORBextractor::DistributeOctTree (Features/ORBextractor.cpp:698-1011) + ExtractorNode::DivideNode
// (:617-676) + the post-distribution fix-up (:1159-1179), one CTA per (level, slot).
//
// The reference keeps a std::list of nodes and push_front()s children; this kernel keeps the same
// information in flat tables and reproduces the list ORDER arithmetically:
//  * every node gets an index in creation (push) order; push_front means the list, front to
//    back, is "alive nodes by DESCENDING index" (root i is stored at index nIni-1-i so that the
//    rule also holds for the push_back'ed roots);
//  * a pass visits the nodes to split in list order (phase 1) or in reverse std::sort order
//    (phase 2), so child indices are a prefix sum over the visiting order;
//  * node boxes depend only on geometry, and a key's child is a pure function of its position
//    and its node's box, so keys are never moved: each key carries its current node id, child
//    populations are counted with atomics, and the per-node "first strict maximum of response in
//    input order" (:984-1007) is an atomicMax over (response, ~emission order) keys;
//  * the tie order of std::sort (unstable, observable -- SURVEY.md 7.3-1) is replayed by one thread
//    with stdsort_emul(); everything else in a pass is CTA-parallel.
#include <cstdlib>

#include "common.cuh"

namespace orbx {

// CTA size QT is a template parameter: small CTAs (64 threads) for the usual quotas of a few hundred keypoints per
// level, so that the one-thread std::sort replay of a CTA overlaps the parallel phases of its SM neighbours; 512
// threads when a level carries thousands of candidates (3840x2160 / 8000 kp), where the candidate loops dominate.


template <int QT>
struct QtS {
  int nNodes, size, gen0, gen1, nS, nProc, nToExpand, cut;
  int part[QT];
};

// (count, UL.x, node id) in one word: std::sort's comparator (:684-696) sees only (count, UL.x)
struct PackedLess {
  __device__ __forceinline__ bool operator()(const u64& a, const u64& b) const { return (a >> 24) < (b >> 24); }
};

// In-place inclusive scan of a[0..n) by the whole CTA.  Returns the total to every thread.  Each thread sums a run of consecutive
// elements; the per-thread sums are scanned with shuffles inside each warp and the warp totals (at most 16) through part[]:
// two CTA barriers per scan.
template <int QT>
__device__ int cta_scan(int* a, int n, int* part) {
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int chunk = (n + QT - 1) / QT;
  const int b = min(tid * chunk, n), e = min(b + chunk, n);
  int s = 0;
  for (int i = b; i < e; ++i) s += a[i];
  int v = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) part[w] = v;
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int i = 0; i < QT / 32; ++i) {
    const int p = part[i];
    if (i < w) base += p;
    total += p;
  }
  int run = base + v - s;
  for (int i = b; i < e; ++i) { run += a[i]; a[i] = run; }
  __syncthreads();
  return total;
}

__device__ __forceinline__ int quadrant(const short4 b, const u32 xy) {
  const int x = (int)(xy & 0xffff), y = (int)(xy >> 16);
  const int midx = b.x + ((b.z - b.x + 1) >> 1);  // UL.x + ceil(w / 2)
  const int midy = b.y + ((b.w - b.y + 1) >> 1);
  return (x >= midx ? 1 : 0) + (y >= midy ? 2 : 0);  // n1, n2, n3, n4
}

// SM (single-frame path, launch_quadtree): the node tables and the candidates' node ids / positions live in shared memory (smNodeCap
// nodes, smCandCap candidates, behind the sort words).  A pass is a handful of dependent steps over those tables -- counts by atomics,
// prefix sums, child boxes, key moves -- and with one frame on the GPU each step's global round trip is the kernel's latency.
template <int QT, bool SM>
__global__ void __launch_bounds__(QT) k_quadtree(const __grid_constant__ Params P, int sortCap, int smNodeCap, int smCandCap) {
  __shared__ QtS<QT> S;
  extern __shared__ __align__(8) u64 sSrt[];  // sortCap packed sort words (launch_quadtree sizes it from the largest quota)
  const int l = blockIdx.x, slot = blockIdx.y, tid = threadIdx.x;
  const LevelInfo& L = P.lv[l];
  const int N = L.quota;
  // test hook (orbx_set_option "qt_tight"): the smallest table the bound above allows, so that the squeeze runs on ordinary scenes
  const int nodeCap = P.optQtTight ? min(L.nodeCap, 5 * N + 4 * L.nIni + 8) : L.nodeCap;
  int* lvlCount = P.lvlCount + (size_t)slot * kMaxLevels + l;
  const int M = min((int)P.candCount[(size_t)slot * kMaxLevels + l], L.candCap);
  if (M == 0 || L.nIni < 1) {
    if (tid == 0) {
      *lvlCount = 0;
      if (L.nIni < 1) atomicMin(P.status + slot, ORBX_ERR_ASPECT);
    }
    return;
  }
  const size_t co = (size_t)slot * P.candTotal + L.candOff;
  const u64* cKey = P.candKey + co;
  const u32* cXY = P.candXY + co;
  u32* cNode = P.candNode + co;
  const size_t no = (size_t)slot * P.nodeTotal + L.nodeOff;
  short4* nBox = P.nodeBox + no;
  int* nCount = P.nodeCount + no;
  int* nSlot = P.nodeSlot + no;
  u64* nBest = P.nodeBest + no;
  int* aux = P.nodeAux + 4 * no;
  int* list = P.nodeList + no;
  int* scan = P.nodeScan + no;
  SortElem* srt = P.nodeSort + no;
  const int* edges = P.rootEdges + L.rootEdgeOff;
  if (SM) {
    u8* sm = reinterpret_cast<u8*>(sSrt + sortCap);
    auto take = [&](size_t bytes) { u8* r = sm; sm += (bytes + 15) & ~(size_t)15; return r; };
    nBest = reinterpret_cast<u64*>(take((size_t)smNodeCap * 8));
    nBox = reinterpret_cast<short4*>(take((size_t)smNodeCap * 8));
    aux = reinterpret_cast<int*>(take((size_t)smNodeCap * 16));
    nCount = reinterpret_cast<int*>(take((size_t)smNodeCap * 4));
    nSlot = reinterpret_cast<int*>(take((size_t)smNodeCap * 4));
    list = reinterpret_cast<int*>(take((size_t)smNodeCap * 4));
    scan = reinterpret_cast<int*>(take((size_t)smNodeCap * 4));
    u32* sNode = reinterpret_cast<u32*>(take((size_t)smCandCap * 4));
    u32* sXY = reinterpret_cast<u32*>(take((size_t)smCandCap * 4));
    if (M <= smCandCap) {  // the usual few thousand candidates; beyond the budget they stay in global memory
      for (int k = tid; k < M; k += QT) { sNode[k] = cNode[k]; sXY[k] = cXY[k]; }
      cNode = sNode; cXY = sXY;
    }
    __syncthreads();
  }

  // roots (:706-768)
  for (int i = tid; i < L.nIni; i += QT) {
    const int r = L.nIni - 1 - i;  // root r stored at index i
    nBox[i] = make_short4((short)edges[r], 0, (short)edges[r + 1], (short)L.distH);
    nCount[i] = 0;
    nSlot[i] = -1;
  }
  __syncthreads();
  for (int k = tid; k < M; k += QT) atomicAdd(nCount + cNode[k], 1);
  __syncthreads();
  if (tid == 0) {
    int s = 0;
    for (int i = 0; i < L.nIni; ++i) s += nCount[i] > 0;
    S.size = s; S.nNodes = L.nIni; S.gen0 = 0; S.gen1 = L.nIni;
  }
  __syncthreads();

  bool phase2 = false;
  for (;;) {
    const int prevSize = S.size, G = S.gen1 - S.gen0;
    int gen0 = S.gen0, gen1 = S.gen1, nNodes = S.nNodes;
    // A pass creates at most 4 children per node of the last generation.  If they might not fit into the node table, the
    // erased nodes are squeezed out first: list order is "alive nodes by descending index", so a renumbering that keeps the
    // relative order changes nothing, and the last generation (all alive) stays the top of the table.  Afterwards
    // nNodes == |L| < N and G <= |L|, so nNodes + 4 G < 5 N <= nodeCap (tables.cu): the table cannot overflow, however
    // slowly a clustered scene lets the list grow.
    if (gen0 > 0 && nNodes + 4 * G > nodeCap) {
      for (int i = tid; i < nNodes; i += QT) scan[i] = nCount[i] > 0;
      __syncthreads();
      const int alive = cta_scan<QT>(scan, nNodes, S.part);
      for (int k = tid; k < M; k += QT) cNode[k] = (u32)(scan[cNode[k]] - 1);
      for (int i = tid; i < nNodes; i += QT) {  // boxes and counts move through the aux table (4 ints per node)
        const int c = nCount[i];
        if (c > 0) {
          const int q = scan[i] - 1;
          const short4 b = nBox[i];
          aux[4 * q] = (int)((u32)(unsigned short)b.x | ((u32)(unsigned short)b.y << 16));
          aux[4 * q + 1] = (int)((u32)(unsigned short)b.z | ((u32)(unsigned short)b.w << 16));
          aux[4 * q + 2] = c;
        }
      }
      __syncthreads();
      for (int q = tid; q < alive; q += QT) {
        const u32 lo = (u32)aux[4 * q], hi = (u32)aux[4 * q + 1];
        nBox[q] = make_short4((short)(lo & 0xffff), (short)(lo >> 16), (short)(hi & 0xffff), (short)(hi >> 16));
        nCount[q] = aux[4 * q + 2];
        nSlot[q] = -1;
      }
      if (tid == 0) { S.nNodes = alive; S.gen1 = alive; S.gen0 = alive - G; }
      __syncthreads();
      nNodes = alive; gen1 = alive; gen0 = alive - G;
    }
    // ---- A. nodes to split, in visiting order ----
    int nS;
    if (!phase2) {
      for (int j = tid; j < G; j += QT) scan[j] = nCount[gen1 - 1 - j] > 1;
      __syncthreads();
      nS = cta_scan<QT>(scan, G, S.part);
      for (int j = tid; j < G; j += QT) {
        const int node = gen1 - 1 - j;
        if (nCount[node] > 1) list[scan[j] - 1] = node;
      }
    } else {
      for (int j = tid; j < G; j += QT) scan[j] = nCount[gen0 + j] > 1;
      __syncthreads();
      nS = cta_scan<QT>(scan, G, S.part);
      // the sort runs in one thread (its tie order is the algorithm's, see stdsort_emul.cuh): keep its
      // elements in shared memory as single 64-bit words, count << 40 | UL.x << 24 | node id
      const bool inSmem = nS <= sortCap;
      for (int j = tid; j < G; j += QT) {
        const int node = gen0 + j, c = nCount[node];
        if (c > 1) {
          if (inSmem) {
            sSrt[scan[j] - 1] = ((u64)(u32)c << 40) | ((u64)(u32)(int)nBox[node].x << 24) | (u64)(u32)node;
          } else {
            SortElem e;
            e.key = ((u64)(u32)c << 32) | (u64)(u32)(int)nBox[node].x;
            e.id = (u32)node; e.pad = 0;
            srt[scan[j] - 1] = e;
          }
        }
      }
      __syncthreads();
      if (tid == 0) {
        if (inSmem) stdsort_emul_t(sSrt, (long)nS, PackedLess());
        else stdsort_emul(srt, nS);
      }
      __syncthreads();
      for (int s = tid; s < nS; s += QT) list[s] = inSmem ? (int)(sSrt[nS - 1 - s] & 0xffffffu) : (int)srt[nS - 1 - s].id;
    }
    __syncthreads();
    if (nS == 0) break;  // nothing divisible: |L| == prevSize
    // ---- B/C. child populations ----
    for (int s = tid; s < nS; s += QT) nSlot[list[s]] = s;
    for (int i = tid; i < 4 * nS; i += QT) aux[i] = 0;
    __syncthreads();
    for (int k = tid; k < M; k += QT) {
      const int nd = (int)cNode[k], s = nSlot[nd];
      if (s >= 0) atomicAdd(aux + 4 * s + quadrant(nBox[nd], cXY[k]), 1);
    }
    __syncthreads();
    // ---- D. child indices = prefix sum over the visiting order; phase 2 stops at |L| >= N ----
    for (int s = tid; s < nS; s += QT)
      scan[s] = (aux[4 * s] > 0) + (aux[4 * s + 1] > 0) + (aux[4 * s + 2] > 0) + (aux[4 * s + 3] > 0);
    if (tid == 0) { S.cut = nS - 1; S.nToExpand = 0; }
    __syncthreads();
    cta_scan<QT>(scan, nS, S.part);
    if (phase2) {
      for (int s = tid; s < nS; s += QT)
        if (prevSize + scan[s] - (s + 1) >= N) atomicMin(&S.cut, s);
      __syncthreads();
    }
    const int nProc = S.cut + 1;
    const int nNew = scan[nProc - 1];
    if (nNodes + nNew > L.nodeCap) {  // unreachable after the squeeze above; kept as a guard
      if (tid == 0) { atomicMin(P.status + slot, ORBX_ERR_CAPACITY); *lvlCount = 0; }
      return;
    }
    // ---- E. create the children (n1..n4 order), erase the parents ----
    for (int s = tid; s < nS; s += QT) {
      const int node = list[s];
      if (s >= nProc) { nSlot[node] = -1; continue; }
      const short4 b = nBox[node];
      const int hx = (b.z - b.x + 1) >> 1, hy = (b.w - b.y + 1) >> 1;
      const int mx = b.x + hx, my = b.y + hy;
      int id = nNodes + (s ? scan[s - 1] : 0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int c = aux[4 * s + q];
        if (c > 0) {
          nBox[id] = make_short4((short)((q & 1) ? mx : b.x), (short)((q & 2) ? my : b.y),
                                 (short)((q & 1) ? b.z : mx), (short)((q & 2) ? b.w : my));
          nCount[id] = c;
          nSlot[id] = -1;
          if (c > 1) atomicAdd(&S.nToExpand, 1);
          aux[4 * s + q] = id++;
        }
      }
      nCount[node] = 0;
    }
    __syncthreads();
    // ---- F. move the keys of the split nodes to their child ----
    for (int k = tid; k < M; k += QT) {
      const int nd = (int)cNode[k], s = nSlot[nd];
      if (s >= 0) cNode[k] = (u32)aux[4 * s + quadrant(nBox[nd], cXY[k])];
    }
    __syncthreads();
    if (tid == 0) {
      S.size = prevSize + nNew - nProc;
      S.gen0 = nNodes; S.gen1 = nNodes + nNew; S.nNodes = nNodes + nNew;
    }
    __syncthreads();
    const int size = S.size;
    if (size >= N || size == prevSize) break;                         // :887-888 / :966-967
    if (!phase2 && size + 3 * S.nToExpand > N) phase2 = true;         // :897
    __syncthreads();
  }
  __syncthreads();
  // ---- best key per node (:984-1007), list order = descending node index, truncation (:1159-1161) ----
  const int nNodes = S.nNodes;
  for (int i = tid; i < nNodes; i += QT) nBest[i] = 0ull;
  __syncthreads();
  for (int k = tid; k < M; k += QT) atomicMax(nBest + cNode[k], cKey[k]);
  for (int j = tid; j < nNodes; j += QT) scan[j] = nCount[nNodes - 1 - j] > 0;
  __syncthreads();
  const int alive = cta_scan<QT>(scan, nNodes, S.part);
  const size_t ko = (size_t)slot * P.kpStride + L.kpOff;
  for (int j = tid; j < nNodes; j += QT) {
    const int node = nNodes - 1 - j;
    if (nCount[node] > 0) {
      const int p = scan[j] - 1;
      if (p < N) {
        const u64 best = nBest[node];
        const u32 order = ~(u32)(best & 0xffffffffull);
        const int cell = (int)(order >> 14), yin = (int)((order >> 7) & 127), xin = (int)(order & 127);
        const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
        // candidate coordinates are relative to (16,16); the reference adds minBorder back (:1171-1173)
        const int x = xin + cj * L.wCell + kMinBorder, y = yin + ci * L.hCell + kMinBorder;
        P.lvlXY[ko + p] = (u32)x | ((u32)y << 16);
        P.lvlResp[ko + p] = (float)(u32)(best >> 32);
      }
    }
  }
  if (tid == 0) *lvlCount = min(alive, N);
}

int launch_quadtree(const Params& P, int nSlots, cudaStream_t st) {
  dim3 grid(P.nlevels, nSlots);
  // a generation has at most ~quota divisible nodes; beyond sortCap the kernel sorts in global memory
  int q = 0, nodeCap = 0, candCap = 0;
  for (int l = 0; l < P.nlevels; ++l) {
    q = std::max(q, P.lv[l].quota); nodeCap = std::max(nodeCap, P.lv[l].nodeCap); candCap = std::max(candCap, P.lv[l].candCap);
  }
  const int sortCap = 2 * q < 256 ? 256 : (2 * q > 4096 ? 4096 : 2 * q);
  const size_t smem = (size_t)sortCap * sizeof(u64);
  // a few frames (the latency path): everything a pass touches in shared memory, if it fits
  candCap = std::min(candCap, 8192);
  const size_t smemSM = smem + (size_t)nodeCap * 48 + (size_t)candCap * 8 + 256;
  if (nSlots <= 8 && q < 1000 && smemSM <= 200 * 1024) {
    k_quadtree<256, true><<<grid, 256, smemSM, st>>>(P, sortCap, nodeCap, candCap);
    return 1;
  }
  // small quotas: 64-thread CTAs give the best throughput when thousands of them share the GPU (whole batches); a chunk of
  // the pipelined host path has few CTAs and waits for the slowest one (level 0), where 256 threads cut the latency by a third
  if (q < 400 && nSlots > 128) k_quadtree<64, false><<<grid, 64, smem, st>>>(P, sortCap, 0, 0);
  else if (q < 1000) k_quadtree<256, false><<<grid, 256, smem, st>>>(P, sortCap, 0, 0);
  else k_quadtree<512, false><<<grid, 512, smem, st>>>(P, sortCap, 0, 0);
  return 1;
}

// Called once per context (after cudaSetDevice): the shared-memory form of k_quadtree opts into its dynamic shared memory.
int configure_quadtree() {
  return cudaFuncSetAttribute(k_quadtree<256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) == cudaSuccess ? ORBX_OK : ORBX_ERR_CUDA;
}

}  // namespace orbx
