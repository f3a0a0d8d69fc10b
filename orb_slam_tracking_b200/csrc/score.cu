// Initializer scoring: the symmetric transfer error of a homography and the epipolar
// distance of a fundamental matrix over the matches of a frame pair,
//   Initializer::CheckHomography   Initialization/Initializer.cpp:268-353
//   Initializer::CheckFundamental  Initialization/Initializer.cpp:355-438
// for the whole hypothesis set of a RANSAC loop at once (FindHomography / FindFundamental, :169-266: 200 hypotheses per
// model and pair), plus the loop's selection rule (:205-209, :257-261).
//
// The score is a float32 running sum over the matches in list order, so its value depends on the order of the
// additions: a hypothesis is therefore walked by ONE thread, match after match, exactly as the reference does, and the
// parallelism comes from the other axes -- one CTA per pair, one thread per hypothesis.  The match list of the pair
// (mvMatches12, Initializer.cpp:27-34: ascending i with vnMatches12[i] >= 0) is compacted in order into shared memory
// once per CTA as (u1, v1, u2, v2); every thread then reads the same entry at the same time (a broadcast, no bank
// conflicts), and the 9 or 18 coefficients of its hypothesis stay in registers.
//
// Arithmetic: float32, in the reference's evaluation order, no FMA contraction (-fmad=false), IEEE division
// (-prec-div=true).  The reference's `1.0 / x` with float x is a double division narrowed to float; a correctly rounded
// float division gives the same value (53 >= 2 * 24 + 2: double rounding is innocuous for division), so the kernel
// uses 1.0f / x and no fp64 instruction.
#include "common.cuh"

namespace orbx {

constexpr int kScoreThreads = 256;

// squared transfer error of (px, py) mapped by the row-major 3x3 M against (qx, qy) (:314-319, :331-336)
__device__ __forceinline__ float transfer_err2(const float (&M)[9], float px, float py, float qx, float qy) {
  const float wInv = 1.0f / (M[6] * px + M[7] * py + M[8]);
  const float mx = (M[0] * px + M[1] * py + M[2]) * wInv;
  const float my = (M[3] * px + M[4] * py + M[5]) * wInv;
  return (qx - mx) * (qx - mx) + (qy - my) * (qy - my);
}
// squared distance of (qx, qy) to the line (l0, l1, l2) (:396-402, :418-424)
__device__ __forceinline__ float line_err2(float l0, float l1, float l2, float qx, float qy) {
  const float num = l0 * qx + l1 * qy + l2;
  return num * num / (l0 * l0 + l1 * l1);
}

// One hypothesis over the list.  The two `score +=` of a match are separate float additions in the reference, so they
// are kept separate here.  A = H21 or F21, B = H12.
template <bool FUND>
__device__ __forceinline__ float walk(const float (&A)[9], const float (&B)[9], const float4* __restrict__ sUV, int N,
                                      float invSigma2, u8* inl) {
  const float limH = (float)5.991, limF = (float)3.841;  // :291 / :372,374
  float total = 0.f;
  for (int i = 0; i < N; ++i) {
    const float4 m = sUV[i];  // u1, v1, u2, v2
    bool ok = true;
    if (!FUND) {
      const float c1 = transfer_err2(B, m.z, m.w, m.x, m.y) * invSigma2;  // frame 2 into frame 1 first
      if (c1 > limH) ok = false; else total += limH - c1;
      const float c2 = transfer_err2(A, m.x, m.y, m.z, m.w) * invSigma2;
      if (c2 > limH) ok = false; else total += limH - c2;
    } else {
      const float c1 = line_err2(A[0] * m.x + A[1] * m.y + A[2], A[3] * m.x + A[4] * m.y + A[5], A[6] * m.x + A[7] * m.y + A[8],
                                 m.z, m.w) * invSigma2;  // F x1 against x2
      if (c1 > limF) ok = false; else total += limH - c1;
      const float c2 = line_err2(A[0] * m.z + A[3] * m.w + A[6], A[1] * m.z + A[4] * m.w + A[7], A[2] * m.z + A[5] * m.w + A[8],
                                 m.x, m.y) * invSigma2;  // F^T x2 against x1
      if (c2 > limF) ok = false; else total += limH - c2;
    }
    if (inl) inl[i] = ok ? 1 : 0;
  }
  return total;
}

template <bool FUND>
__global__ void __launch_bounds__(kScoreThreads) k_score_models(const ScoreArgs S) {
  extern __shared__ __align__(16) u8 smemRaw[];
  float4* sUV = reinterpret_cast<float4*>(smemRaw);                    // [stride]
  float* sScore = reinterpret_cast<float*>(smemRaw + (size_t)S.stride * sizeof(float4));  // [nHyp]
  __shared__ int sWarp[kScoreThreads / 32];
  __shared__ int sBase, sBest;

  const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s1 = pair * S.pairStep, s2 = s1 + 1;
  const int n1 = min(S.n[s1], S.stride), n2 = S.n[s2];
  const orbx_keypoint* k1 = S.keysUn + (size_t)s1 * S.stride;
  const orbx_keypoint* k2 = S.keysUn + (size_t)s2 * S.stride;
  const int* m12 = S.matches12 + (size_t)pair * S.stride;

  // ordered compaction of the match list
  if (tid == 0) sBase = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n1; i0 += kScoreThreads) {
    const int i = i0 + tid;
    const int j = i < n1 ? m12[i] : -1;
    const bool on = j >= 0 && j < n2;
    const u32 bal = __ballot_sync(0xffffffffu, on);
    if (lane == 0) sWarp[warp] = __popc(bal);
    __syncthreads();
    int pos = sBase + __popc(bal & ((1u << lane) - 1u));
    for (int w = 0; w < warp; ++w) pos += sWarp[w];
    if (on) sUV[pos] = make_float4(k1[i].x, k1[i].y, k2[j].x, k2[j].y);
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < kScoreThreads / 32; ++w) t += sWarp[w];
      sBase += t;
    }
    __syncthreads();
  }
  const int N = sBase;
  if (tid == 0) S.nList[pair] = N;

  // one thread per hypothesis
  for (int h = tid; h < S.nHyp; h += kScoreThreads) {
    float A[9], B[9];
    const float* pa = S.hypA + ((size_t)pair * S.nHyp + h) * 9;
#pragma unroll
    for (int k = 0; k < 9; ++k) A[k] = pa[k];
    if (!FUND) {
      const float* pb = S.hypB + ((size_t)pair * S.nHyp + h) * 9;
#pragma unroll
      for (int k = 0; k < 9; ++k) B[k] = pb[k];
    } else {
#pragma unroll
      for (int k = 0; k < 9; ++k) B[k] = 0.f;
    }
    u8* inl = S.inlAll ? S.inlAll + ((size_t)pair * S.nHyp + h) * N : nullptr;
    const float sc = walk<FUND>(A, B, sUV, N, S.invSigmaSquare, inl);
    S.scores[(size_t)pair * S.nHyp + h] = sc;
    sScore[h] = sc;
  }
  __syncthreads();
  // the RANSAC selection: the running best starts at 0 and only a strictly larger score replaces it
  if (tid == 0) {
    float score = 0.f;
    int best = -1;
    for (int it = 0; it < S.nHyp; ++it)
      if (sScore[it] > score) { score = sScore[it]; best = it; }
    sBest = best;
    S.best[pair] = best;
  }
  if (!S.inlBest) return;
  __syncthreads();
  const int best = sBest;
  u8* out = S.inlBest + (size_t)pair * S.stride;
  if (best < 0) {
    for (int i = tid; i < S.stride; i += kScoreThreads) out[i] = 0;
    return;
  }
  float A[9], B[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    A[k] = S.hypA[((size_t)pair * S.nHyp + best) * 9 + k];
    B[k] = FUND ? 0.f : S.hypB[((size_t)pair * S.nHyp + best) * 9 + k];
  }
  for (int i = tid; i < S.stride; i += kScoreThreads) {
    u8 f = 0;
    if (i < N) walk<FUND>(A, B, sUV + i, 1, S.invSigmaSquare, &f);  // the flags do not depend on the running sum
    out[i] = f;
  }
}

int launch_score_models(const ScoreArgs& S, int nPairs, bool fundamental, cudaStream_t st) {
  const size_t smem = (size_t)S.stride * sizeof(float4) + (size_t)S.nHyp * sizeof(float);
  if (fundamental) {
    cudaFuncSetAttribute(k_score_models<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    k_score_models<true><<<nPairs, kScoreThreads, smem, st>>>(S);
  } else {
    cudaFuncSetAttribute(k_score_models<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    k_score_models<false><<<nPairs, kScoreThreads, smem, st>>>(S);
  }
  return 1;
}

}  // namespace orbx
