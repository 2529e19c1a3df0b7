/* srt_detmath.h -- the numeric decree of the hot path (host + device, header only).
 *
 * GLSL leaves the precision of sin/cos/pow/normalize/length and the contraction of a*b+c to the
 * implementation (GLSL 4.50 section 4.7.1), and the shader's random numbers are a chaotic function of
 * every upstream ulp (raytrace_utils.glsl:28-30,44-54).  So "the reference's result" is only defined
 * once those choices are pinned.  This header pins them, once, for BOTH the CPU oracle (oracle/) and
 * the CUDA kernels (simple-ray-tracer_b200/csrc/): every function here is built from IEEE-754
 * correctly rounded + - * / sqrt fma, float<->int conversions and bit moves only, so g++ and nvcc give
 * identical bits provided neither contracts on its own (g++ -ffp-contract=off, nvcc -fmad=false) and
 * neither flushes denormals (no -ffast-math, no -use_fast_math / -ftz).
 *
 * Decrees:
 *   min(x,y) = (y<x)?y:x ; max(x,y) = (x<y)?y:x                 (GLSL 4.50 section 8.3, NaN behaviour included)
 *   dot / cross / mat*vec     contracted left to right with fma  (what a GPU GLSL compiler emits)
 *   length(v)=sqrt(dot(v,v)) ; normalize(v)=v/length(v) (IEEE divide) ; reflect(I,N)=I-2*dot(N,I)*N
 *   mix(a,b,t)=a*(1-t)+b*t ; fract(x)=x-floor(x) ; round = half away from zero ; pow(x,5)=x*x*x*x*x
 *   sin/cos : 3-constant Cody-Waite reduction by pi/2 + degree-7/8 minimax polynomials
 *   pow(x,y), y fractional (sRGB curve only): exp2(y*log2(x)) from the polynomials below
 *   int(x) : truncation; NaN -> 0; saturating at the int32 range
 */
#ifndef SRT_DETMATH_H_
#define SRT_DETMATH_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define SRT_HD __host__ __device__ __forceinline__
#else
#define SRT_HD static inline
#endif

namespace srt {

struct V3 {
  float x, y, z;
};

SRT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
SRT_HD V3 v3(float s) { return v3(s, s, s); }
SRT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
SRT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
SRT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
SRT_HD V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
SRT_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
SRT_HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
SRT_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
SRT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
SRT_HD V3 operator-(float s, V3 a) { return v3(s - a.x, s - a.y, s - a.z); }

SRT_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
SRT_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}

SRT_HD float fma_(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return fmaf(a, b, c);
#endif
}

SRT_HD float inf_() { return u2f(0x7f800000u); }
SRT_HD bool isnan_(float x) { return x != x; }
SRT_HD bool isinf_(float x) { return (f2u(x) & 0x7fffffffu) == 0x7f800000u; }

/* GLSL 4.50 section 8.3: min(x,y) returns y if y<x, otherwise x (so min(NaN,y)=NaN, min(x,NaN)=x). */
SRT_HD float min_(float x, float y) { return (y < x) ? y : x; }
SRT_HD float max_(float x, float y) { return (x < y) ? y : x; }
SRT_HD float clamp_(float x, float lo, float hi) { return min_(max_(x, lo), hi); }
SRT_HD float saturate_(float x) { return clamp_(x, 0.0f, 1.0f); }
SRT_HD float abs_(float x) { return u2f(f2u(x) & 0x7fffffffu); }
SRT_HD V3 min_(V3 a, V3 b) { return v3(min_(a.x, b.x), min_(a.y, b.y), min_(a.z, b.z)); }
SRT_HD V3 max_(V3 a, V3 b) { return v3(max_(a.x, b.x), max_(a.y, b.y), max_(a.z, b.z)); }
SRT_HD V3 abs_(V3 a) { return v3(abs_(a.x), abs_(a.y), abs_(a.z)); }

SRT_HD float dot2_(float ax, float ay, float bx, float by) { return fma_(ay, by, ax * bx); }
SRT_HD float dot_(V3 a, V3 b) { return fma_(a.z, b.z, fma_(a.y, b.y, a.x * b.x)); }
SRT_HD V3 cross_(V3 a, V3 b) {
  return v3(fma_(a.y, b.z, -(b.y * a.z)),
            fma_(a.z, b.x, -(b.z * a.x)),
            fma_(a.x, b.y, -(b.x * a.y)));
}
SRT_HD float sqrt_(float x) {
#if defined(__CUDA_ARCH__)
  return __fsqrt_rn(x);
#else
  return sqrtf(x);
#endif
}
SRT_HD float div_(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdiv_rn(a, b);
#else
  return a / b;
#endif
}
/* a / b where a is often +-0 (a normal's component on an axis-aligned face, a colour channel of a saturated material,
 * N.L of a light behind the surface).  The GPU's correctly rounded division refuses a zero numerator on its fast path and
 * calls a 116-instruction subroutine (13 % of the shading kernel's instructions on the Rubik scene,
 * profiles/r02n_logic_by_line.txt); IEEE says what comes out without dividing: 0 / b = +-0 (sign = xor) for any b that is
 * not 0 or NaN, and 0 / 0 = NaN.  Same value as div_(a, b) on both sides (every NaN counts as the same value). */
SRT_HD float div0_(float a, float b) {
#if defined(__CUDA_ARCH__)
  const uint32_t ub = f2u(b) & 0x7fffffffu;
  const bool zero = (a == 0.0f) && (ub <= 0x7f800000u);
  const float q = __fdiv_rn(zero ? 1.0f : a, zero ? 1.0f : b);  /* (a zero numerator never reaches the divider, nor does 0 / 0) */
  const float z = ub == 0u ? u2f(0x7fffffffu) : u2f((f2u(a) ^ f2u(b)) & 0x80000000u);
  return zero ? z : q;
#else
  return a / b;
#endif
}
// 1/x, correctly rounded: the same bits as div_(1.0f, x) (IEEE division is correctly rounded too), fewer instructions
SRT_HD float rcp_(float x) {
#if defined(__CUDA_ARCH__)
  return __frcp_rn(x);
#else
  return 1.0f / x;
#endif
}
SRT_HD float length_(V3 a) { return sqrt_(dot_(a, a)); }
SRT_HD V3 normalize_(V3 a) { float l = length_(a); return v3(div0_(a.x, l), div0_(a.y, l), div0_(a.z, l)); }
SRT_HD V3 div0_(V3 a, float s) { return v3(div0_(a.x, s), div0_(a.y, s), div0_(a.z, s)); }
SRT_HD V3 reflect_(V3 I, V3 N) { float k = 2.0f * dot_(N, I); return I - k * N; }
SRT_HD float mix_(float a, float b, float t) { return a * (1.0f - t) + b * t; }
SRT_HD V3 mix_(V3 a, V3 b, float t) { return v3(mix_(a.x, b.x, t), mix_(a.y, b.y, t), mix_(a.z, b.z, t)); }
SRT_HD float floor_(float x) { return floorf(x); }
SRT_HD float fract_(float x) { return x - floor_(x); }
SRT_HD float round_(float x) { return roundf(x); }
SRT_HD float pow5_(float x) { return x * x * x * x * x; }

/* int(x) of GLSL (undefined for NaN / out of range there; pinned here). */
SRT_HD int32_t f2i_(float x) {
  if (isnan_(x)) return 0;
  if (x >= 2147483648.0f) return 2147483647;
  if (x <= -2147483648.0f) return (-2147483647 - 1);
  return (int32_t)x;
}

/* column-major mat4 (m[c*4+r]) times (v, w); returns xyz.  raytrace_compute.glsl:146-147. */
SRT_HD V3 mat4_mul_(const float* m, V3 v, float w) {
  V3 r;
  r.x = fma_(m[12], w, fma_(m[8], v.z, fma_(m[4], v.y, m[0] * v.x)));
  r.y = fma_(m[13], w, fma_(m[9], v.z, fma_(m[5], v.y, m[1] * v.x)));
  r.z = fma_(m[14], w, fma_(m[10], v.z, fma_(m[6], v.y, m[2] * v.x)));
  return r;
}

/* ---- sin / cos ---------------------------------------------------------------------------- */
SRT_HD float sincos_core_(float x, int want_cos) {
  if (!(abs_(x) <= 1.0e8f)) {
    if (isnan_(x) || isinf_(x)) return u2f(0x7fc00000u);
    x = fmodf(x, 6.2831854820251465f); /* fmod is exact in IEEE arithmetic on both sides */
  }
  float j = rintf(x * 0.63661977236758134f); /* nearest multiple of pi/2, ties to even */
  int32_t q = (int32_t)j;
  float r = fma_(-j, 1.5703125f, x);
  r = fma_(-j, 4.837512969970703125e-4f, r);
  r = fma_(-j, 7.54978995489188216e-8f, r);
  float z = r * r;
  /* sin(r) and cos(r) on [-pi/4, pi/4] */
  float ps = fma_(fma_(-1.9515295891e-4f, z, 8.3321608736e-3f), z, -1.6666654611e-1f);
  float s = fma_(ps * z, r, r);
  float pc = fma_(fma_(2.443315711809948e-5f, z, -1.388731625493765e-3f), z, 4.166664568298827e-2f);
  float c = fma_(pc * z, z, fma_(-0.5f, z, 1.0f));
  q = (q + want_cos) & 3;
  float v = (q & 1) ? c : s;
  return (q & 2) ? -v : v;
}
SRT_HD float sin_(float x) { return sincos_core_(x, 0); }
SRT_HD float cos_(float x) { return sincos_core_(x, 1); }

/* ---- pow for the sRGB curve (raytrace_utils.glsl:179) --------------------------------------- */
SRT_HD float log2_(float x) {
  if (isnan_(x) || x < 0.0f) return u2f(0x7fc00000u);
  if (x == 0.0f) return -inf_();
  if (isinf_(x)) return x;
  float bias = 0.0f;
  if (x < 1.17549435e-38f) { x = x * 8388608.0f; bias = -23.0f; }
  uint32_t ix = f2u(x);
  int32_t d = (int32_t)ix - 0x3f3504f3;
  /* floor(d / 2^23) written without an implementation-defined shift of a negative value */
  int32_t e = (d >= 0) ? (d >> 23) : -(((-d) + 0x7fffff) >> 23);
  float m = u2f((uint32_t)((int32_t)ix - e * 8388608)); /* m in [sqrt(1/2), sqrt(2)) */
  float f = m - 1.0f;
  float s = div_(f, 2.0f + f);
  float z = s * s;
  float p = fma_(fma_(fma_(fma_(0.2222222222f, z, 0.2857142857f), z, 0.4f), z, 0.6666666667f), z, 2.0f);
  float ln_m = s * p;
  return fma_(ln_m, 1.4426950408889634f, (float)e + bias);
}
SRT_HD float exp2_(float y) {
  if (isnan_(y)) return y;
  if (y >= 128.0f) return inf_();
  if (y < -150.0f) return 0.0f;
  float n = rintf(y);
  float f = y - n; /* [-0.5, 0.5] */
  float p = 1.535336188319500e-4f;
  p = fma_(p, f, 1.339887440266574e-3f);
  p = fma_(p, f, 9.618437357674640e-3f);
  p = fma_(p, f, 5.550332471162809e-2f);
  p = fma_(p, f, 2.402264791363012e-1f);
  p = fma_(p, f, 6.931472028550421e-1f);
  p = fma_(p, f, 1.0f);
  int32_t ni = (int32_t)n;
  /* scale by 2^ni in two exact steps so denormal results round once */
  int32_t h = ni / 2;
  float s1 = u2f((uint32_t)(h + 127) << 23);
  float s2 = u2f((uint32_t)(ni - h + 127) << 23);
  return p * s1 * s2;
}
SRT_HD float pow_(float x, float y) {
  if (x == 0.0f) return (y > 0.0f) ? 0.0f : ((y == 0.0f) ? 1.0f : inf_());
  return exp2_(y * log2_(x));
}

}  // namespace srt

#endif /* SRT_DETMATH_H_ */
