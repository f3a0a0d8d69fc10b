// Orientation, steered BRIEF and output packing.
//   IC_Angle               Features/ORBextractor.cpp:103-159   (on the UNBLURRED level, :1184-1188)
//   computeOrbDescriptor   Features/ORBextractor.cpp:169-228   (on the blurred level, :1613)
//   output packing         Features/ORBextractor.cpp:1558-1652 (level-major, pt *= scale, lapping split)
// Moments are integer sums (order-independent, so a shuffle reduction is
// exact); cv::fastAtan2 is replayed in float32 with explicit round-to-nearest multiplies and adds
// (this file is also compiled with -fmad=false); the rotated sample coordinates use the same
// separate float multiply/add roundings and round-half-even as cvRound on x86-64.
#include "common.cuh"

namespace orbx {

__constant__ signed char c_pattern[256 * 4] = {
#include "orb_pattern.inc"
};

// cv::fastAtan2 (degrees), SURVEY.md appendix A4
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float scale = (float)(180.0 / 3.1415926535897932384626433832795);
  const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale;
  const float p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
  const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

// cos / sin of the keypoint angle.  The reference's `cos(angle)` / `sin(angle)` (:174) take a float under <cmath> and
// `using namespace std`, i.e. glibc's cosf / sinf.  glibc is no part of the reference tree; its algorithm (glibc >= 2.28,
// sysdeps/ieee754/flt-32/s_sincosf.h: one multiply by 2/pi * 2^24 gives the quadrant, the reduced argument and two short
// polynomials run in double, ONE rounding to float) is restated here with explicit fused multiply-adds where the x86-64
// __cosf_fma / __sinf_fma builds contract them.  The same restatement in the oracle equals the libm of the test machine on
// every float of [0, 2 pi] (tools/probe/glibc_sincosf_check.c), and this one equals the oracle's (tests/test_trig_port.py).
__device__ __forceinline__ void glibc_sincosf(float y, float& sn, float& cs) {
  const u32 top = (__float_as_uint(y) >> 20) & 0x7ffu;
  if (top >= 0x42fu) { sn = sinf(y); cs = cosf(y); return; }  // |y| >= 120, inf, nan: outside the restated range (an angle is <= 2 pi)
  if (top < 0x398u) { sn = y; cs = 1.0f; return; }            // |y| < 2^-12
  double x = (double)y;
  int n = 0;
  if (top >= 0x3f4u) {                                        // |y| >= pi / 4
    n = (__double2int_rz(__dmul_rn(x, 0x1.45F306DC9C883p+23)) + 0x800000) >> 24;
    x = __fma_rn(-(double)n, 0x1.921FB54442D18p0, x);
  }
  const double x2 = __dmul_rn(x, x);
  const double xs = ((n + 1) & 2) ? -x : x;                   // sign of the sine in quadrants 0..3: + - - +
  const double x3 = __dmul_rn(xs, x2), s1 = __fma_rn(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7), x7 = __dmul_rn(x3, x2);
  const float sp = __double2float_rn(__fma_rn(x7, s1, __fma_rn(x3, -0x1.555545995a603p-3, xs)));
  // second table (n & 2): the cosine polynomial with every coefficient negated = the negated result (rounding is symmetric)
  const double x4 = __dmul_rn(x2, x2), c2 = __fma_rn(x2, 0x1.99343027bf8c3p-16, -0x1.6c087e89a359dp-10);
  const double c1 = __fma_rn(x2, -0x1.ffffffd0c621cp-2, 1.0), x6 = __dmul_rn(x4, x2);
  const float cq = __double2float_rn(__fma_rn(x6, c2, __fma_rn(x4, 0x1.55553e1068f19p-5, c1)));
  const float cp = (n & 2) ? -cq : cq;
  if (n & 1) { sn = cp; cs = sp; } else { sn = sp; cs = cp; }
}

constexpr int kWarpsPerCta = 8;
constexpr int kKpPerCta = 32;           // keypoint slots per CTA: 4 per warp, one per thread of warp 0 for the trigonometry
constexpr int kRowBytes = 4 * kPatchWords;                   // 64: a staged patch row
constexpr int kDiscBytes = kDiscRows * kRowBytes;            // 1984: TMA box of the unblurred level (rows y - 15 .. y + 15)
constexpr int kPatchBytes = kPatchRows * kRowBytes;          // 2368: TMA box of the blurred level (rows y - 18 .. y + 18)
constexpr int kPatchBuf = (kPatchBytes + 127) & ~127;        // ring buffers stay 128-byte aligned
#ifndef ORBX_DESC_CTAS
#define ORBX_DESC_CTAS 5
#endif

// Three phases per CTA of 32 keypoint slots:
//   1. IC_Angle, one warp per keypoint: the 31 x 31 disc is read as aligned 32-bit words (lanes over (row, word)),
//      each byte weighted by its column offset u and masked by |u| <= umax[|v|].
//   2. fastAtan2 and cos/sin, one LANE per keypoint of the warp (the double-precision libm calls are the longest
//      instruction stream of the kernel: the warp runs them once for its up to 4 keypoints; no CTA barrier -- after the
//      set-up a warp never waits for another).
//   3. steered BRIEF, one warp per keypoint: lane = descriptor byte gathers its 16 samples from the blurred patch.
// Both phases read the keypoint's neighbourhood from shared memory, where ONE TMA box copy per keypoint and image put it
// (cp.async.bulk.tensor.3d, 16 words x 31 or 37 rows, issued by one lane into the warp's two-deep ring, completion on the
// buffer's mbarrier): the kernel is bound by L1 wavefronts and load latency, and the box copies use neither the LSU pipe
// nor registers.  A warp's copies form one sequence -- the unblurred discs of its (up to 4) keypoints, then the blurred
// patches -- and copy i + 2 is issued as soon as buffer i has been consumed, so the blurred patches of its first two
// keypoints are in flight while the CTA is at the barriers around phase 2.
__global__ void __launch_bounds__(kWarpsPerCta * 32, ORBX_DESC_CTAS) k_describe(const __grid_constant__ Params P, const __grid_constant__ PatchMaps M) {
  __shared__ __align__(128) u32 sPatch[kWarpsPerCta][2][kPatchBuf / 4];
  __shared__ float4 sPat[256];  // pattern pair (x0, y0, x1, y1) as floats, [i * 32 + byte]: conflict-free for lane = byte
  __shared__ __align__(8) u64 sBar[kWarpsPerCta][2];
  __shared__ u32 sDD[32];       // per disc row: 0x80808080 | umax[|v|] * 0x01010101 (byte-wise column limit of IC_Angle)
  __shared__ u32 sXY[kKpPerCta];
  __shared__ signed char sLevel[kKpPerCta];  // -1: empty slot
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  {
    const int b = tid & 31, i = tid >> 5;  // 256 threads: pair (b, i) -> pattern row 8b + i
    const signed char* p = c_pattern + 4 * (8 * b + i);
    sPat[i * 32 + b] = make_float4((float)p[0], (float)p[1], (float)p[2], (float)p[3]);
    if (tid < 31) sDD[tid] = 0x80808080u | ((u32)P.umax[tid < kHalfPatch ? kHalfPatch - tid : tid - kHalfPatch] * 0x01010101u);
  }
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sBar[warp][0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&sBar[warp][1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  const int slot = blockIdx.y;
  const int gBase = blockIdx.x * kKpPerCta;
  if (tid < kKpPerCta) {
    const int g = gBase + tid;
    int l = -1;
    if (g < P.kpStride) {
      // the slot's coordinates are fetched whether or not it is occupied: the load then runs beside the count's
      // instead of behind it (the other seven warps wait for this one at the barrier)
      const u32 xy = __ldg(P.lvlXY + (size_t)slot * P.kpStride + g);
      l = 0;
      while (l + 1 < P.nlevels && g >= P.lv[l + 1].kpOff) ++l;
      if (g - P.lv[l].kpOff >= P.lvlCount[(size_t)slot * kMaxLevels + l]) l = -1;
      sXY[tid] = xy;
    }
    sLevel[tid] = (signed char)l;
  }
  __syncthreads();

  // the warp's occupied keypoint slots (k = warp + 8 j), one per byte of kk
  u32 kk = 0;
  int nv = 0;
#pragma unroll
  for (int j = 0; j < kKpPerCta / kWarpsPerCta; ++j) {
    const int k = warp + kWarpsPerCta * j;
    if (sLevel[k] >= 0) { kk |= (u32)k << (8 * nv); ++nv; }
  }
  // copy i of the warp's sequence into buffer i & 1: the disc of keypoint i (i < nv) or the blurred patch of keypoint i - nv
  auto issue = [&](int i) {
    if (lane == 0) {
      const bool blurred = i >= nv;
      const int k = (int)(kk >> (8 * (blurred ? i - nv : i))) & 31;
      const int l = sLevel[k];
      const u32 xy = sXY[k];
      const int x = (int)(xy & 0xffff), y = (int)(xy >> 16);
      const u32 bar = smem_addr(&sBar[warp][i & 1]);
      const CUtensorMap* map = blurred ? &M.blur[l] : &M.pyr[l];
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(blurred ? kPatchBytes : kDiscBytes) : "memory");
      asm volatile(
          "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
          ::"r"(smem_addr(&sPatch[warp][i & 1][0])), "l"(reinterpret_cast<u64>(map)), "r"(((x - 18) & ~15) >> 2), "r"(blurred ? y - 18 : y - kHalfPatch),
            "r"(slot + P.slot0), "r"(bar)
          : "memory");
    }
  };
  auto wait = [&](int i) {
    const u32 bar = smem_addr(&sBar[warp][i & 1]), parity = (u32)(i >> 1) & 1u;
    u32 done = 0;
    while (!done) {
      asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                   : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    }
  };
  if (nv > 0) { issue(0); issue(1); }  // 2 nv >= 2 copies

  // ---- phase 1: IC_Angle ----
  // lanes 0..26 = 3 rows x 9 words; the disc's columns [x - 15, x + 15] lie in 9 aligned words of the staged rows.  A lane
  // keeps its word column, so its four column offsets u_j are fixed and only the row's half-width d = umax[|v|] changes:
  // byte j counts iff |u_j| <= d
  int k10 = 0, k01 = 0;  // lane i: the moments of the warp's keypoint i
  for (int i = 0; i < nv; ++i) {
    const int k = (int)(kk >> (8 * i)) & 31;
    const int x = (int)(sXY[k] & 0xffff);
    const int xa = (x - 18) & ~15;                 // first staged column
    const int w0 = (x - kHalfPatch - xa) >> 2;     // first of the disc's 9 words in a staged row
    const int lr = lane / 9, w = lane - 9 * lr;
    const int u0 = xa + 4 * (w0 + w) - x;
    // |u_j| per byte (< 128); byte j of a row counts iff umax >= |u_j|: bit 7 of (0x80 | umax) - |u_j| survives
    const u32 apk = (u32)abs(u0) | ((u32)abs(u0 + 1) << 8) | ((u32)abs(u0 + 2) << 16) | ((u32)abs(u0 + 3) << 24);
    const int upk = (u0 & 255) | (((u0 + 1) & 255) << 8) | (((u0 + 2) & 255) << 16) | (((u0 + 3) & 255) << 24);  // s8 x 4
    const u32* const pw = &sPatch[warp][i & 1][lr * kPatchWords + w0 + w];
    int m10 = 0, m01 = 0;
    wait(i);
    if (lane < 27) {
      // m01 = sum over rows of (r - 15) * rowsum with r = rb + lr: T = sum of the bytes, U = sum of rb * bytes (rb is a
      // compile-time byte), m01 = U + (lr - 15) * T
      u32 T = 0, U = 0;
#pragma unroll
      for (int rb = 0; rb < 33; rb += 3) {
        if (rb + lr < kDiscRows) {
          // bit 7 of each byte of (0x80 | umax) - |u| says "inside the disc"; the byte permute in sign-replication mode widens it to 0xff / 0x00
          u32 inside;  // (__byte_perm ignores the replication bit of the selector nibbles: PTX prmt spelled out)
          asm("prmt.b32 %0, %1, 0, 0xba98;" : "=r"(inside) : "r"(sDD[rb + lr] - apk));
          const u32 word = pw[rb * kPatchWords] & inside;
          asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(m10) : "r"(word), "r"(upk), "r"(m10));
          T = __dp4a(word, 0x01010101u, T);
          U = __dp4a(word, (u32)rb * 0x01010101u, U);
        }
      }
      m01 = (int)U + (lr - kHalfPatch) * (int)T;
    }
    __syncwarp();  // every lane has read buffer i & 1
    if (i + 2 < 2 * nv) issue(i + 2);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    if (lane == i) { k10 = m10; k01 = m01; }
  }

  // ---- phase 2: one lane per keypoint of the warp ----
  float myCos = 0.f, mySin = 0.f;
  if (lane < nv) {
    const float angle = fast_atan2_deg((float)k01, (float)k10);
    const float a = __fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f));
    if (P.trigMode == 2) glibc_sincosf(a, mySin, myCos);
    else if (P.trigMode == 0) { myCos = (float)cos((double)a); mySin = (float)sin((double)a); }
    else { myCos = cosf(a); mySin = sinf(a); }
    P.lvlAngle[(size_t)slot * P.kpStride + gBase + ((kk >> (8 * lane)) & 31)] = angle;
  }
  __syncwarp();

  // ---- phase 3: steered BRIEF, lane = descriptor byte ----
  for (int i = 0; i < nv; ++i) {
    const int k = (int)(kk >> (8 * i)) & 31;
    const int x = (int)(sXY[k] & 0xffff);
    const int xa = (x - 18) & ~15;
    const u8* centre = reinterpret_cast<const u8*>(&sPatch[warp][(nv + i) & 1][0]) + 18 * kRowBytes + (x - xa);
    const float co = __shfl_sync(0xffffffffu, myCos, i), si = __shfl_sync(0xffffffffu, mySin, i);
    wait(nv + i);
    u32 val = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 q = sPat[j * 32 + lane];
      const float x0 = q.x, y0 = q.y, x1 = q.z, y1 = q.w;
      const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, si), __fmul_rn(y0, co)));
      const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, co), __fmul_rn(y0, si)));
      const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, si), __fmul_rn(y1, co)));
      const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, co), __fmul_rn(y1, si)));
      const int t0 = centre[r0 * kRowBytes + c0], t1 = centre[r1 * kRowBytes + c1];
      val |= (u32)(t0 < t1) << j;
    }
    __syncwarp();  // every lane has read buffer (nv + i) & 1
    if (nv + i + 2 < 2 * nv) issue(nv + i + 2);
    P.lvlDesc[((size_t)slot * P.kpStride + gBase + k) * 32 + lane] = (u8)val;
  }
}

// ---- output packing: one CTA per slot ----
// CTA size PK: 256 threads for whole batches, 1024 for a few frames (a keypoint per thread: the kernel is a chain of dependent
// loads per keypoint, and a single frame's kernel time is its latency)
template <int PK>
__global__ void __launch_bounds__(PK) k_pack(const __grid_constant__ Params P, int lap0, int lap1) {
  __shared__ int sPart[PK / 32];
  __shared__ int sOff[kMaxLevels + 1];
  const int slot = blockIdx.x, tid = threadIdx.x;
  if (tid == 0) {
    int s = 0;
    for (int l = 0; l < P.nlevels; ++l) { sOff[l] = s; s += P.lvlCount[(size_t)slot * kMaxLevels + l]; }
    sOff[P.nlevels] = s;
  }
  __syncthreads();
  const int total = sOff[P.nlevels];
  const int chunk = (total + PK - 1) / PK;
  const int b = min(tid * chunk, total), e = min(b + chunk, total);
  const size_t base = (size_t)slot * P.kpStride;
  const float fl0 = (float)lap0, fl1 = (float)lap1;
  // pass 1: lapping keypoints in my chunk
  int nl = 0;
  {
    int l = 0;
    for (int i = b; i < e; ++i) {
      while (i >= sOff[l + 1]) ++l;
      const u32 xy = P.lvlXY[base + P.lv[l].kpOff + (i - sOff[l])];
      float x = (float)(xy & 0xffff);
      if (l != 0) x = __fmul_rn(x, P.lv[l].scale);
      nl += (x >= fl0 && x <= fl1);
    }
  }
  int nLap;
  const int lapIncl = cta_prefix_sum<PK>(nl, sPart, &nLap);
  int lapBefore = lapIncl - nl;
  // pass 2: write
  {
    int l = 0;
    for (int i = b; i < e; ++i) {
      while (i >= sOff[l + 1]) ++l;
      const size_t src = base + P.lv[l].kpOff + (i - sOff[l]);
      const u32 xy = P.lvlXY[src];
      float x = (float)(xy & 0xffff), y = (float)(xy >> 16);
      if (l != 0) { x = __fmul_rn(x, P.lv[l].scale); y = __fmul_rn(y, P.lv[l].scale); }
      const bool lap = (x >= fl0 && x <= fl1);
      const int dst = lap ? total - 1 - lapBefore : i - lapBefore;
      lapBefore += lap;
      orbx_keypoint k;
      k.x = x; k.y = y; k.size = P.lv[l].patchSize; k.angle = P.lvlAngle[src]; k.response = P.lvlResp[src];
      k.octave = l; k.class_id = -1;
      P.outKp[base + dst] = k;
      const uint4* d = reinterpret_cast<const uint4*>(P.lvlDesc + src * 32);
      uint4* o = reinterpret_cast<uint4*>(P.outDesc + (base + dst) * 32);
      o[0] = d[0]; o[1] = d[1];
    }
  }
  if (tid == 0) { P.outN[slot] = total; P.outMono[slot] = total - nLap; }
}

// the restated cosf / sinf on n floats (orbx_sincosf_probe: test hook for tests/test_trig_port.py)
__global__ void k_sincosf_probe(const float* __restrict__ y, long long n, float* __restrict__ sn, float* __restrict__ cs) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) glibc_sincosf(y[i], sn[i], cs[i]);
}
void launch_sincosf_probe(const float* y, long long n, float* sn, float* cs, cudaStream_t st) {
  k_sincosf_probe<<<1184, 256, 0, st>>>(y, n, sn, cs);
}

int launch_describe(const Params& P, const PatchMaps& M, int nSlots, int lap0, int lap1, cudaStream_t st) {
  dim3 grid((P.kpStride + kKpPerCta - 1) / kKpPerCta, nSlots);
  k_describe<<<grid, kWarpsPerCta * 32, 0, st>>>(P, M);
  if (nSlots <= 8) k_pack<1024><<<nSlots, 1024, 0, st>>>(P, lap0, lap1);
  else k_pack<256><<<nSlots, 256, 0, st>>>(P, lap0, lap1);
  return 2;
}

}  // namespace orbx
