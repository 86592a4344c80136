// Test infrastructure (oracle side): host-side exhaustive pin of the smoothing kernels' constant division.
// Build and run: gcc -O2 -mfma -fopenmp -ffp-contract=off oracle/check_div9.c -o /tmp/check_div9 -lm && /tmp/check_div9
// exhaustive check of the branch-free s/9: q0 = s*y; r = fma(-9,q0,s); q1 = fma(r,y,q0); result = (r ordered-!= 0) ? q1 : q0
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <omp.h>
static inline float f(uint32_t u){ float x; memcpy(&x,&u,4); return x; }
static inline uint32_t b(float x){ uint32_t u; memcpy(&u,&x,4); return u; }
int main(){
  const float y = 0.111111112f;
  unsigned long long bad = 0; uint32_t first = 0xffffffffu;
  #pragma omp parallel for reduction(+:bad) schedule(static)
  for (long long i = 0; i < (1ll<<32); ++i) {
    volatile float s = f((uint32_t)i);
    float q0 = s * y;
    float r = fmaf(-9.0f, q0, s);
    float q1 = fmaf(r, y, q0);
    float got = (r < 0.f || r > 0.f) ? q1 : q0;
    float want = s / 9.0f;
    int same = (b(got) == b(want)) || (got != got && want != want);
    if (!same) { bad++; 
      #pragma omp critical
      { if ((uint32_t)i < first) first = (uint32_t)i; } }
  }
  printf("mismatches %llu first %08x\n", bad, first);
  // jpeg.cu::div255: k / 255 for k = 0..255 by the same correction, y = RN(1/255)
  int bad255 = 0;
  const float y255 = 0.00392156886f;
  for (int k = 0; k < 256; ++k) {
    const float kf = (float)k, q0 = kf * y255, r = fmaf(-255.0f, q0, kf), q = fmaf(r, y255, q0);
    if (b(q) != b(kf / 255.0f)) bad255++;
  }
  printf("k/255 mismatches %d\n", bad255);
  return bad != 0 || bad255 != 0;
}
