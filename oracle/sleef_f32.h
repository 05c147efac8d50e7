/*
 * sleef_f32.h - TEST INFRASTRUCTURE (part of the oracle, see vmas_oracle.h): scalar restatement of four single-precision
 * functions of SLEEF, the vector math library PyTorch's CPU Vectorized<float> is built on.
 *
 * Third-party arithmetic on the hot path: the reference's torch.logaddexp (core.py:2820) evaluates
 * exp and log1p through ATen's Vectorized<float>::exp() / ::log1p(), which call Sleef_expf{8,16}_u10 / Sleef_log1pf{8,16}_u10
 * (SLEEF 3.x, Boost Software License; bundled in torch 2.11.0's libtorch_cpu.so, source not under /root/reference).
 * These are the published algorithms of SLEEF's sleefsimdsp.c (xexpf, xlog1pf, xsinf_u1, xcosf_u1) in their FMA form
 * (the AVX2 / AVX-512 builds torch dispatches to), with SLEEF's double-float helpers of df.h.  Every operation is one
 * rounded fp32 instruction or one fmaf, so the results are bit-identical to the vector code: tests/test_oracle_sleef.py
 * checks that against torch itself, and tools/sleef_check.c against the Sleef_* symbols exported by libtorch_cpu.so.
 *
 * sinf / cosf are here for reference only: torch.sin / torch.cos / torch.exp as UNARY ops on the CPU do not use SLEEF but
 * Intel MKL's VML (vsSin / vsCos / vsExp, closed source), see vmas_oracle.c `tg`.
 */
#ifndef VMAS_ORACLE_SLEEF_F32_H_
#define VMAS_ORACLE_SLEEF_F32_H_
#include <math.h>
#include <stdint.h>
#include <string.h>
typedef struct { float x, y; } sl_f2;
static inline float sl_fma(float a, float b, float c) { return fmaf(a, b, c); }
static inline sl_f2 sl_f2mk(float x, float y) { sl_f2 r = {x, y}; return r; }
static inline sl_f2 dfadd2_f_f(float x, float y) { float s = x + y; float v = s - x; return sl_f2mk(s, (x - (s - v)) + (y - v)); }
static inline sl_f2 dfadd2_f2_f(sl_f2 x, float y) { float s = x.x + y; float v = s - x.x; float t = (x.x - (s - v)) + (y - v); return sl_f2mk(s, t + x.y); }
static inline sl_f2 dfadd_f2_f(sl_f2 x, float y) { float s = x.x + y; return sl_f2mk(s, ((x.x - s) + y) + x.y); }
static inline sl_f2 dfadd_f_f2(float x, sl_f2 y) { float s = x + y.x; return sl_f2mk(s, ((x - s) + y.x) + y.y); }
static inline sl_f2 dfadd_f_f(float x, float y) { float s = x + y; return sl_f2mk(s, (x - s) + y); }
static inline sl_f2 dfadd_f2_f2(sl_f2 x, sl_f2 y) { float s = x.x + y.x; return sl_f2mk(s, (((x.x - s) + y.x) + x.y) + y.y); }
static inline sl_f2 dfsqu(sl_f2 x) { float s = x.x * x.x; return sl_f2mk(s, sl_fma(x.x + x.x, x.y, sl_fma(x.x, x.x, -s))); }
static inline sl_f2 dfmul_f2_f2(sl_f2 x, sl_f2 y) { float s = x.x * y.x; return sl_f2mk(s, sl_fma(x.x, y.y, sl_fma(x.y, y.x, sl_fma(x.x, y.x, -s)))); }
static inline float dfmul_f_f2_f2(sl_f2 x, sl_f2 y) { return sl_fma(x.x, y.x, sl_fma(x.y, y.x, x.x * y.y)); }
static inline sl_f2 dfmul_f2_f(sl_f2 x, float y) { float s = x.x * y; return sl_f2mk(s, sl_fma(x.y, y, sl_fma(x.x, y, -s))); }
static inline sl_f2 dfdiv(sl_f2 n, sl_f2 d) {
  float t = 1.0f / d.x; float s = n.x * t; float u = sl_fma(t, n.x, -s);
  float v = sl_fma(-d.y, t, sl_fma(-d.x, t, 1.0f));
  return sl_f2mk(s, sl_fma(s, v, sl_fma(n.y, t, u)));
}
#define SL_PI_A2f 3.1414794921875f
#define SL_PI_B2f 0.00011315941810607910156f
#define SL_PI_C2f 1.9841872589410058936e-09f
static inline float sl_negif(float u, int neg) { uint32_t b; memcpy(&b, &u, 4); if (neg) b ^= 0x80000000u; memcpy(&u, &b, 4); return u; }
/* |d| < 125 only */
static inline float sleef_sinf_u10(float d) {
  float u = rintf(d * (float)M_1_PI);
  int q = (int)u;
  float v = sl_fma(u, -SL_PI_A2f, d);
  sl_f2 s = dfadd2_f_f(v, u * -SL_PI_B2f);
  s = dfadd_f2_f(s, u * -SL_PI_C2f);
  sl_f2 t = s;
  s = dfsqu(s);
  u = 2.6083159809786593541503e-06f;
  u = sl_fma(u, s.x, -0.0001981069071916863322258f);
  u = sl_fma(u, s.x, 0.00833307858556509017944336f);
  sl_f2 x = dfadd_f_f2(1.0f, dfmul_f2_f2(dfadd_f_f(-0.166666597127914428710938f, u * s.x), s));
  u = dfmul_f_f2_f2(t, x);
  u = sl_negif(u, (q & 1) == 1);
  if (d == 0.0f && signbit(d)) u = d;
  return u;
}
static inline float sleef_cosf_u10(float d) {
  float dq = sl_fma(rintf(sl_fma(d, (float)M_1_PI, -0.5f)), 2.0f, 1.0f);
  int q = (int)dq;
  sl_f2 s = dfadd2_f_f(d, dq * (-SL_PI_A2f * 0.5f));
  s = dfadd2_f2_f(s, dq * (-SL_PI_B2f * 0.5f));
  s = dfadd2_f2_f(s, dq * (-SL_PI_C2f * 0.5f));
  sl_f2 t = s;
  s = dfsqu(s);
  float u = 2.6083159809786593541503e-06f;
  u = sl_fma(u, s.x, -0.0001981069071916863322258f);
  u = sl_fma(u, s.x, 0.00833307858556509017944336f);
  sl_f2 x = dfadd_f_f2(1.0f, dfmul_f2_f2(dfadd_f_f(-0.166666597127914428710938f, u * s.x), s));
  u = dfmul_f_f2_f2(t, x);
  u = sl_negif(u, (q & 2) == 0);
  return u;
}
static inline float sl_pow2if(int q) { uint32_t b = (uint32_t)(q + 0x7f) << 23; float f; memcpy(&f, &b, 4); return f; }
static inline float sl_ldexp2(float d, int e) { return d * sl_pow2if(e >> 1) * sl_pow2if(e - (e >> 1)); }
static inline float sleef_expf_u10(float d) {
  int q = (int)rintf(d * 1.442695040888963407359924681001892137426645954152985934135449406931f);
  float s = sl_fma((float)q, -0.693145751953125f, d);
  s = sl_fma((float)q, -1.428606765330187045e-06f, s);
  float u = 0.000198527617612853646278381f;
  u = sl_fma(u, s, 0.00139304355252534151077271f);
  u = sl_fma(u, s, 0.00833336077630519866943359f);
  u = sl_fma(u, s, 0.0416664853692054748535156f);
  u = sl_fma(u, s, 0.166666671633720397949219f);
  u = sl_fma(u, s, 0.5f);
  u = 1.0f + sl_fma(s * s, u, s);
  u = sl_ldexp2(u, q);
  if (d < -104.0f) u = 0.0f;
  if (100.0f < d) u = INFINITY;
  return u;
}
static inline int sl_ilogb2k(float d) { uint32_t b; memcpy(&b, &d, 4); return (int)((b >> 23) & 0xff) - 0x7f; }
static inline float sl_ldexp3(float d, int e) { uint32_t b; memcpy(&b, &d, 4); b += (uint32_t)e << 23; memcpy(&d, &b, 4); return d; }
/* AVX512 variant uses vgetexp; the AVX2 variant ilogb2k: same value for normal inputs */
static inline float sleef_log1pf_u10(float d) {
  float dp1 = d + 1.0f;
  int o = dp1 < 1.17549435e-38f;
  if (o) dp1 = dp1 * (4294967296.0f * 4294967296.0f);
  int e = sl_ilogb2k(dp1 * (1.0f / 0.75f));
  float t = sl_ldexp3(1.0f, -e);
  float m = sl_fma(d, t, t - 1.0f);
  if (o) e -= 64;
  sl_f2 s = dfmul_f2_f(sl_f2mk(0.69314718246459960938f, -1.904654323148236017e-09f), (float)e);
  sl_f2 x = dfdiv(sl_f2mk(m, 0.0f), dfadd_f_f(2.0f, m));
  float x2 = x.x * x.x;
  t = +0.3027294874e+0f;
  t = sl_fma(t, x2, +0.3996108174e+0f);
  t = sl_fma(t, x2, +0.6666694880e+0f);
  s = dfadd_f2_f2(s, sl_f2mk(x.x * 2.0f, x.y * 2.0f));
  s = dfadd_f2_f(s, x2 * x.x * t);
  float r = s.x + s.y;
  if (d > 1e+38f) r = INFINITY;
  if (-1.0f > d) r = NAN;
  if (d == -1.0f) r = -INFINITY;
  if (d == 0.0f && signbit(d)) r = -0.0f;
  return r;
}
#endif
