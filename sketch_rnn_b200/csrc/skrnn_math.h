// skrnn_math.h -- deterministic float32 math shared by device code and by host checkers.
//
// The sampler (Model.sample, reference model.py:187-264) must produce bit-identical stroke
// sequences on the GPU and in the CPU oracle when both are fed the same uniform draws.  libm
// and the CUDA math library round transcendental functions differently, and compilers fuse
// a*b+c differently, so every operation on that path goes through this header: explicit
// round-to-nearest add / mul / fma / div / sqrt (never contracted), and polynomial
// exp / log / tanh / sincos built only from those.  Compile host users with
// -ffp-contract=off.  Polynomial coefficients are the classic Cephes single-precision ones.
#ifndef SKRNN_MATH_H_
#define SKRNN_MATH_H_

#include <stdint.h>
#include <string.h>
#include <math.h>

#if defined(__CUDA_ARCH__)
#define SKM_FN __device__ __forceinline__
#define SKM_ADD(a, b) __fadd_rn((a), (b))
#define SKM_SUB(a, b) __fsub_rn((a), (b))
#define SKM_MUL(a, b) __fmul_rn((a), (b))
#define SKM_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define SKM_DIV(a, b) __fdiv_rn((a), (b))
#define SKM_SQRT(a) __fsqrt_rn((a))
#define SKM_DADD(a, b) __dadd_rn((a), (b))
#else
#if defined(__CUDACC__)
#define SKM_FN __host__ __device__ inline
#else
#define SKM_FN static inline
#endif
#define SKM_ADD(a, b) ((float)((float)(a) + (float)(b)))
#define SKM_SUB(a, b) ((float)((float)(a) - (float)(b)))
#define SKM_MUL(a, b) ((float)((float)(a) * (float)(b)))
#define SKM_FMA(a, b, c) fmaf((a), (b), (c))
#define SKM_DIV(a, b) ((float)((float)(a) / (float)(b)))
#define SKM_SQRT(a) sqrtf((a))
#define SKM_DADD(a, b) ((double)((double)(a) + (double)(b)))
#endif

SKM_FN uint32_t skm_f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
#endif
}
SKM_FN float skm_u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// exp(x), |rel err| ~ 1e-7.  Input clamped to [-87, 88] (so exp(-inf) returns ~1.6e-38, not 0).
SKM_FN float skm_exp(float x) {
  if (!(x > -87.0f)) x = -87.0f;  // also catches NaN -> smallest
  if (x > 88.0f) x = 88.0f;
  float n = floorf(SKM_FMA(x, 1.44269504088896341f, 0.5f));
  float r = SKM_FMA(n, -0.693359375f, x);
  r = SKM_FMA(n, 2.12194440e-4f, r);
  float p = 1.9875691500e-4f;
  p = SKM_FMA(p, r, 1.3981999507e-3f);
  p = SKM_FMA(p, r, 8.3334519073e-3f);
  p = SKM_FMA(p, r, 4.1665795894e-2f);
  p = SKM_FMA(p, r, 1.6666665459e-1f);
  p = SKM_FMA(p, r, 5.0000001201e-1f);
  float rr = SKM_MUL(r, r);
  float y = SKM_ADD(SKM_FMA(p, rr, r), 1.0f);
  int32_t e = (int32_t)n + 127;
  return SKM_MUL(y, skm_u2f((uint32_t)e << 23));
}

// log(x) for x >= 0.  x below FLT_MIN (zero, denormal) returns -inf.
SKM_FN float skm_log(float x) {
  uint32_t u = skm_f2u(x);
  if (x < 1.17549435e-38f) return skm_u2f(0xff800000u);
  int32_t e = (int32_t)(u >> 23) - 127;
  float m = skm_u2f((u & 0x007fffffu) | 0x3f800000u);  // [1, 2)
  if (m > 1.41421356237f) {
    m = SKM_MUL(m, 0.5f);
    e += 1;
  }
  float f = SKM_SUB(m, 1.0f);
  float s = SKM_DIV(f, SKM_ADD(2.0f, f));
  float z = SKM_MUL(s, s);
  // 2*atanh(s) = 2s (1 + z/3 + z^2/5 + z^3/7 + z^4/9 + z^5/11)
  float p = 0.0909090909f;
  p = SKM_FMA(p, z, 0.1111111111f);
  p = SKM_FMA(p, z, 0.1428571429f);
  p = SKM_FMA(p, z, 0.2f);
  p = SKM_FMA(p, z, 0.3333333333f);
  p = SKM_MUL(p, z);
  float two_s = SKM_ADD(s, s);
  float lm = SKM_FMA(two_s, p, two_s);
  float fe = (float)e;
  return SKM_FMA(fe, 0.693359375f, SKM_FMA(fe, -2.12194440e-4f, lm));
}

SKM_FN float skm_tanh(float x) {
  float ax = fabsf(x);
  float r;
  if (ax < 0.625f) {
    float z = SKM_MUL(x, x);
    float p = -5.70498872745e-3f;
    p = SKM_FMA(p, z, 2.06390887954e-2f);
    p = SKM_FMA(p, z, -5.37397155531e-2f);
    p = SKM_FMA(p, z, 1.33314422036e-1f);
    p = SKM_FMA(p, z, -3.33332819422e-1f);
    return SKM_FMA(SKM_MUL(p, z), x, x);
  }
  if (ax >= 9.0f) {
    r = 1.0f;
  } else {
    float e = skm_exp(SKM_ADD(ax, ax));
    r = SKM_SUB(1.0f, SKM_DIV(2.0f, SKM_ADD(e, 1.0f)));
  }
  return x < 0.0f ? -r : r;
}

SKM_FN float skm_sigmoid(float x) { return SKM_DIV(1.0f, SKM_ADD(1.0f, skm_exp(-x))); }

// (sin, cos)(2*pi*u) for u in [0, 1).
SKM_FN void skm_sincos2pi(float u, float *s_out, float *c_out) {
  float a = SKM_MUL(u, 4.0f);
  float kf = floorf(SKM_ADD(a, 0.5f));
  float g = SKM_SUB(a, kf);  // [-0.5, 0.5)
  float t = SKM_MUL(g, 1.57079632679489662f);
  float z = SKM_MUL(t, t);
  float ps = -1.9515295891e-4f;
  ps = SKM_FMA(ps, z, 8.3321608736e-3f);
  ps = SKM_FMA(ps, z, -1.6666654611e-1f);
  float s = SKM_FMA(SKM_MUL(ps, z), t, t);
  float pc = 2.443315711809948e-5f;
  pc = SKM_FMA(pc, z, -1.388731625493765e-3f);
  pc = SKM_FMA(pc, z, 4.166664568298827e-2f);
  float c = SKM_ADD(SKM_FMA(SKM_MUL(pc, z), z, SKM_MUL(-0.5f, z)), 1.0f);
  int k = ((int)kf) & 3;
  float so = s, co = c;
  if (k == 1) { so = c; co = -s; }
  else if (k == 2) { so = -s; co = -c; }
  else if (k == 3) { so = -c; co = s; }
  *s_out = so;
  *c_out = co;
}

// Box-Muller: two standard normals from two uniforms in [0,1).  u1 is mapped to (0,1].
SKM_FN void skm_box_muller(float u1, float u2, float *z1, float *z2) {
  float l = skm_log(SKM_SUB(1.0f, u1));
  float r = SKM_SQRT(SKM_MUL(-2.0f, l));
  float s, c;
  skm_sincos2pi(u2, &s, &c);
  *z1 = SKM_MUL(r, c);
  *z2 = SKM_MUL(r, s);
}

// Philox4x32-10 (Salmon et al. 2011), counter-based RNG for in-kernel dropout masks and
// sampler uniforms.  Pure integer arithmetic: identical on host and device by construction.
SKM_FN void skm_philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                           uint32_t k1, uint32_t out[4]) {
  for (int i = 0; i < 10; ++i) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// Counter hash for in-kernel dropout masks: three rounds of the lowbias32 integer finaliser over
// (seed, unit, timestep, global row).  Integer only, so host and device agree.  The row is mixed in last: a
// thread that owns one hidden unit computes the (unit, t) prefix once per step and pays one round per row.
SKM_FN uint32_t skm_mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
SKM_FN uint32_t skm_drop_unit(uint64_t seed, uint32_t unit) { return skm_mix32(unit * 0xC2B2AE3Du ^ (uint32_t)seed); }
SKM_FN uint32_t skm_drop_prefix(uint64_t seed, uint32_t unit_hash, uint32_t t) {
  return skm_mix32(unit_hash ^ (t * 0x85EBCA77u) ^ (uint32_t)(seed >> 32));
}
SKM_FN uint32_t skm_drop_finish(uint32_t prefix, uint32_t row) { return skm_mix32(prefix ^ (row * 0x9E3779B1u)); }
SKM_FN uint32_t skm_drop_bits(uint64_t seed, uint32_t row, uint32_t t, uint32_t unit) {
  return skm_drop_finish(skm_drop_prefix(seed, skm_drop_unit(seed, unit), t), row);
}
// 24-bit uniform in [0,1): exactly representable, same on host and device.
SKM_FN float skm_u01(uint32_t r) { return (float)(r >> 8) * 5.9604644775390625e-8f; }

#endif  // SKRNN_MATH_H_
