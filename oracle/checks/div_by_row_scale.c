/* Test infrastructure (never linked into the product): exhaustive proof that the per-token quantisers may replace
 * the IEEE division  q = a / s  (reference: csrc/quantization/fp8/common.cu:74, `x / scale`) by
 *     rc = RN(1 / s);  q0 = RN(a * rc);  r = fma(-s, q0, a);  q1 = fma(r, rc, q0)
 * with rc computed once per token.  The quotient's bits depend only on the two significands (scaling by powers of
 * two is exact while nothing over- or underflows), so all 2^23 significands of s against every significand `a` can
 * have are enumerated: a is a value that was rounded to bf16 (8 significant bits) or f16 (11 bits) before the
 * division, which is what every caller divides.  Prints the number of mismatches; 0 is the pass.
 *   gcc -O2 -mfma -fopenmp -o /tmp/div_check oracle/checks/div_by_row_scale.c -lm && /tmp/div_check 11 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline float bits2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int main(int argc, char** argv) {
  const int abits = argc > 1 ? atoi(argv[1]) : 11;   /* significant bits of a, including the hidden one */
  const uint32_t na = 1u << (abits - 1);
  long long bad = 0;
#pragma omp parallel for reduction(+ : bad) schedule(dynamic, 4096)
  for (uint32_t ms = 0; ms < (1u << 23); ++ms) {
    const float s = bits2f(0x3f800000u | ms);         /* [1, 2) */
    const float rc = 1.0f / s;
    for (uint32_t ma = 0; ma < na; ++ma) {
      const float a = bits2f(0x3f800000u | (ma << (24 - abits)));
      const float want = a / s;
      const float q0 = a * rc;
      const float r = fmaf(-s, q0, a);
      const float q1 = fmaf(r, rc, q0);
      if (f2bits(want) != f2bits(q1)) ++bad;
    }
  }
  printf("a significand bits %d: %lld mismatches over %llu quotients\n", abits, bad,
         (unsigned long long)na << 23);
  return bad != 0;
}
