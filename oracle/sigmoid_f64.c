/* C restatement of the build's sigmoid contract (oracle/moe_oracle.py:sigmoid_f64, csrc/router.cu:sigmoid_f64).
 * TEST INFRASTRUCTURE ONLY.  The contract replaces the reference's `gating_output.sigmoid()`
 * (test_biased_grouped_topk.py:22) by a fixed sequence of IEEE-754 double operations, so that every implementation
 * returns the same bits.  tests/test_oracle.py compiles this file (gcc -O2 -ffp-contract=off) and holds the torch
 * restatement bit-equal to it on every bf16 input and on random f32 inputs: two independent implementations of
 * the sequence agreeing is the evidence that a third one (the CUDA kernel) can.
 *   gcc -O2 -ffp-contract=off -shared -fPIC -o oracle/_build/libsigmoid_f64.so oracle/sigmoid_f64.c -lm */
#include <math.h>
#include <stdint.h>
#include <string.h>

static const double C[14] = {0x1.0000000000000p+0,  0x1.0000000000000p+0,  0x1.0000000000000p-1,  0x1.5555555555555p-3,
                             0x1.5555555555555p-5,  0x1.1111111111111p-7,  0x1.6c16c16c16c17p-10, 0x1.a01a01a01a01ap-13,
                             0x1.a01a01a01a01ap-16, 0x1.71de3a556c734p-19, 0x1.27e4fb7789f5cp-22, 0x1.ae64567f544e4p-26,
                             0x1.1eed8eff8d898p-29, 0x1.6124613a86d09p-33};

float sglb_sigmoid_f64(float x) {
  volatile double t = -(double)x;
  if (t != t) return x;
  t = t < -50.0 ? -50.0 : (t > 120.0 ? 120.0 : t);
  const double k = rint(t * 0x1.71547652b82fep+0); /* default rounding mode: half to even */
  const double r = (t - k * 0x1.62e42fee00000p-1) - k * 0x1.a39ef35793c76p-33;
  double p = C[13];
  for (int n = 12; n >= 0; --n) p = p * r + C[n];
  const int64_t bits = ((int64_t)k + 1023) << 52;
  double two_k;
  memcpy(&two_k, &bits, 8);
  return (float)(1.0 / (1.0 + p * two_k));
}

void sglb_sigmoid_f64_array(const float* x, float* out, long n) {
  for (long i = 0; i < n; ++i) out[i] = sglb_sigmoid_f64(x[i]);
}
