// Exhaustive check of a restatement of glibc's sinf / cosf (sysdeps/ieee754/flt-32/s_sincosf.h, s_sinf.c, s_cosf.c; glibc >= 2.28,
// the ARM optimized-routines algorithm) against the libm of this machine on EVERY float in [0, hi] (default hi = 6.2831855f * 1.0001:
// the keypoint angles of computeOrbDescriptor, Features/ORBextractor.cpp:173-174, are kpt.angle * (pi / 180) with kpt.angle in [0, 360)).
// Two variants: every a * b + c contracted into a fused multiply-add (what the ifunc-selected __sinf_fma / __cosf_fma of an x86-64
// glibc run on any CPU with FMA) and none contracted (the generic build).  Prints the number of mismatching floats for each.
//   gcc -O2 -ffp-contract=off -pthread -o glibc_sincosf_check glibc_sincosf_check.c -lm && ./glibc_sincosf_check
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static const double kHpiInv = 0x1.45F306DC9C883p+23, kHpi = 0x1.921FB54442D18p0;
static const double kC[5] = {0x1p0, -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16};
static const double kS[3] = {-0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13};

static inline uint32_t asu(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline uint32_t top12(float f) { return (asu(f) >> 20) & 0x7ff; }
#define MADD(F, a, b, c) ((F) ? fma((a), (b), (c)) : (a) * (b) + (c))

// sinf_poly: quadrant n even -> sine polynomial, odd -> cosine polynomial (negated when neg)
static inline float poly(int F, double x, double x2, int n, int neg) {
  if ((n & 1) == 0) {
    const double x3 = x * x2, s1 = MADD(F, x2, kS[2], kS[1]), x7 = x3 * x2, s = MADD(F, x3, kS[0], x);
    return (float)MADD(F, x7, s1, s);
  }
  const double sg = neg ? -1.0 : 1.0;
  const double x4 = x2 * x2, c2 = MADD(F, x2, sg * kC[4], sg * kC[3]), c1 = MADD(F, x2, sg * kC[1], sg * kC[0]), x6 = x4 * x2;
  const double c = MADD(F, x4, sg * kC[2], c1);
  return (float)MADD(F, x6, c2, c);
}
static inline float port(int F, float y, int isCos) {
  double x = y;
  if (top12(y) < top12(0x1.921FB6p-1f)) {
    if (top12(y) < top12(0x1p-12f)) return isCos ? 1.0f : y;
    return poly(F, x, x * x, isCos, 0);
  }
  if (top12(y) < top12(120.0f)) {
    const double r = x * kHpiInv;
    const int n = ((int32_t)r + 0x800000) >> 24;
    x = F ? fma(-(double)n, kHpi, x) : x - n * kHpi;
    const double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    return poly(F, x * s, x * x, n ^ isCos, n & 2);
  }
  return isCos ? cosf(y) : sinf(y);  // large arguments: not restated (never reached by an angle in degrees)
}

typedef struct { uint32_t lo, hi; unsigned long bad[4]; uint32_t first[4]; } Job;
static void* run(void* p) {
  Job* j = (Job*)p;
  for (uint32_t u = j->lo; u < j->hi; ++u) {
    float y; memcpy(&y, &u, 4);
    const float c = cosf(y), s = sinf(y);
    for (int F = 0; F < 2; ++F) {
      if (asu(port(F, y, 1)) != asu(c)) { if (!j->bad[2 * F]++) j->first[2 * F] = u; }
      if (asu(port(F, y, 0)) != asu(s)) { if (!j->bad[2 * F + 1]++) j->first[2 * F + 1] = u; }
    }
  }
  return 0;
}
int main(int argc, char** argv) {
  const float hi = argc > 1 ? (float)atof(argv[1]) : 6.2831855f * 1.0001f;
  const int nt = argc > 2 ? atoi(argv[2]) : 8;
  const uint32_t top = asu(hi) + 1;
  pthread_t th[64]; Job jobs[64];
  for (int t = 0; t < nt; ++t) {
    memset(&jobs[t], 0, sizeof(Job));
    jobs[t].lo = (uint32_t)((uint64_t)top * t / nt); jobs[t].hi = (uint32_t)((uint64_t)top * (t + 1) / nt);
    pthread_create(&th[t], 0, run, &jobs[t]);
  }
  unsigned long bad[4] = {0, 0, 0, 0}; uint32_t first[4] = {0, 0, 0, 0};
  for (int t = 0; t < nt; ++t) { pthread_join(th[t], 0); for (int k = 0; k < 4; ++k) { if (jobs[t].bad[k] && !bad[k]) first[k] = jobs[t].first[k]; bad[k] += jobs[t].bad[k]; } }
  printf("floats checked: %u (0 .. %.9g)\n", top, hi);
  printf("no contraction : cosf mismatches %lu (first 0x%08x), sinf mismatches %lu (first 0x%08x)\n", bad[0], first[0], bad[1], first[1]);
  printf("all fused (fma): cosf mismatches %lu (first 0x%08x), sinf mismatches %lu (first 0x%08x)\n", bad[2], first[2], bad[3], first[3]);
  return 0;
}
