// csrc/rt_math.cuh -- device-side FP32/FP64 vector math with the reference's exact
// rounding sequence (C/v3math.c:56-192 and the typed expressions of C/raycast.c).
// Every float product/sum is a separately rounded __fmul_rn/__fadd_rn (the reference is
// built without FMA contraction, C/Makefile:4); the only fused operations are the ones
// DESIGN.md proves equivalent (the sphere reject test).  IEEE division and square root
// (nvcc defaults -prec-div=true -prec-sqrt=true -ftz=false; never --use_fast_math).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RT_DEV __device__ __forceinline__

struct v3 { float x, y, z; };

RT_DEV float mul32(float a, float b) { return __fmul_rn(a, b); }
RT_DEV float add32(float a, float b) { return __fadd_rn(a, b); }
RT_DEV float sub32(float a, float b) { return __fsub_rn(a, b); }
RT_DEV double mul64(double a, double b) { return __dmul_rn(a, b); }
RT_DEV double add64(double a, double b) { return __dadd_rn(a, b); }
RT_DEV double sub64(double a, double b) { return __dsub_rn(a, b); }

RT_DEV v3 mk3(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_DEV v3 sub3(v3 a, v3 b) { return mk3(sub32(a.x, b.x), sub32(a.y, b.y), sub32(a.z, b.z)); }

// v3_dot_product, C/v3math.c:69-71: (a0*b0 + a1*b1) + a2*b2, three products, two sums.
RT_DEV float dot3(v3 a, v3 b) {
  return add32(add32(mul32(a.x, b.x), mul32(a.y, b.y)), mul32(a.z, b.z));
}
// The double sum d0^2 + d1^2 + d2^2 (squares of floats are exact in double).
RT_DEV double sumsq_d(v3 a) {
  double x = (double)a.x, y = (double)a.y, z = (double)a.z;
  return add64(add64(mul64(x, x), mul64(y, y)), mul64(z, z));
}
// v3_length, C/v3math.c:169-172.
RT_DEV float len3(v3 a) { return (float)sqrt(sumsq_d(a)); }
// v3_normalize, C/v3math.c:180-192: unchanged on zero length, else three divisions.
RT_DEV v3 norm3(v3 a) {
  float l = len3(a);
  if (l == 0.0f) return a;
  return mk3(__fdiv_rn(a.x, l), __fdiv_rn(a.y, l), __fdiv_rn(a.z, l));
}
RT_DEV v3 norm3_len(v3 a, float &l) {
  l = len3(a);
  if (l == 0.0f) return a;
  return mk3(__fdiv_rn(a.x, l), __fdiv_rn(a.y, l), __fdiv_rn(a.z, l));
}
// v3_reflect, C/v3math.c:144-160: v - n*(2*dot(v,n)).
RT_DEV v3 reflect3(v3 v, v3 n) {
  float s = mul32(2.0f, dot3(v, n));
  return mk3(sub32(v.x, mul32(n.x, s)), sub32(v.y, mul32(n.y, s)), sub32(v.z, mul32(n.z, s)));
}
// "a" of the sphere quadratic, C/raycast.c:583 (ray invariant).
RT_DEV float ray_a(v3 d) { return (float)sumsq_d(d); }

// x^20 by the same multiplication chain as oracle/rt_oracle.c pow20().
RT_DEV double pow20(double x) {
  double x2 = mul64(x, x), x4 = mul64(x2, x2), x5 = mul64(x4, x), x10 = mul64(x5, x5);
  return mul64(x10, x10);
}
// alpha^a0 as oracle/rt_oracle.c pow_a0(): integer exponents 0..64 by binary powering.
RT_DEV double pow_a0(double x, float a0) {
  if (a0 >= 0.0f && a0 <= 64.0f && (float)(int)a0 == a0) {
    int e = (int)a0;
    double r = 1.0, b = x;
    while (e) { if (e & 1) r = mul64(r, b); b = mul64(b, b); e >>= 1; }
    return r;
  }
  return pow(x, (double)a0);
}

// ---- literal intersections (candidate / slow paths) -------------------------------
// ray_intersect_sphere, C/raycast.c:576-600, typed as the reference types it.
// Returns false when the discriminant is negative; t may be <= 0 / inf / NaN.
__device__ __noinline__ bool sphere_t_literal(v3 o, v3 d, float4 s, float &t) {
  v3 tmp = sub3(o, mk3(s.x, s.y, s.z));
  float a = ray_a(d);
  float b = mul32(2.0f, dot3(d, tmp));
  float c = (float)sub64((double)dot3(tmp, tmp), mul64((double)s.w, (double)s.w));
  float disc = (float)sub64(mul64((double)b, (double)b), (double)mul32(mul32(4.0f, a), c));
  if (disc < 0.0f) return false;
  double sq = sqrt((double)disc), den = mul64(2.0, (double)a), nb = -(double)b;
  t = (float)__ddiv_rn(sub64(nb, sq), den);
  if (t < 0.0f) t = (float)__ddiv_rn(add64(nb, sq), den);
  return true;
}

// ray_intersect_plane, C/raycast.c:545-562. num = dot(o - pos, n) is passed in
// (origin-only, shared by every ray leaving the same point).
RT_DEV bool plane_t(float num, v3 d, v3 n, float &t) {
  float den = dot3(d, n);
  if (den == 0.0f) return false;
  t = __fdiv_rn(mul32(-1.0f, num), den);
  return !(t < 0.0f);
}

// Quadric coefficients a..j.
struct quad10 { float A, B, C, D, E, F, G, H, I, J; };

// c_q of C/raycast.c:629-638 (origin-only).
RT_DEV double quadric_cq(v3 o, const quad10 &q) {
  double ox = o.x, oy = o.y, oz = o.z;
  double s = mul64((double)q.A, mul64(ox, ox));
  s = add64(s, mul64((double)q.B, mul64(oy, oy)));
  s = add64(s, mul64((double)q.C, mul64(oz, oz)));
  s = add64(s, (double)mul32(mul32(q.D, o.x), o.y));
  s = add64(s, (double)mul32(mul32(q.E, o.x), o.z));
  s = add64(s, (double)mul32(mul32(q.F, o.y), o.z));
  s = add64(s, (double)mul32(q.G, o.x));
  s = add64(s, (double)mul32(q.H, o.y));
  s = add64(s, (double)mul32(q.I, o.z));
  s = add64(s, (double)q.J);
  return s;
}

// ray_intersect_quadric, C/raycast.c:614-656, with c_q (as float) precomputed.
RT_DEV bool quadric_t(v3 o, v3 d, const quad10 &q, float c_q, float &t) {
  double dx = d.x, dy = d.y, dz = d.z, ox = o.x, oy = o.y, oz = o.z;
  double s = mul64((double)q.A, mul64(dx, dx));
  s = add64(s, mul64((double)q.B, mul64(dy, dy)));
  s = add64(s, mul64((double)q.C, mul64(dz, dz)));
  s = add64(s, (double)mul32(mul32(q.D, d.x), d.y));
  s = add64(s, (double)mul32(mul32(q.E, d.x), d.z));
  s = add64(s, (double)mul32(mul32(q.F, d.y), d.z));
  float a_q = (float)s;
  double b = mul64(mul64(mul64(2.0, (double)q.A), ox), dx);
  b = add64(b, mul64(mul64(mul64(2.0, (double)q.B), oy), dy));
  b = add64(b, mul64(mul64(mul64(2.0, (double)q.C), oz), dz));
  b = add64(b, (double)mul32(q.D, add32(mul32(o.x, d.y), mul32(o.y, d.x))));
  b = add64(b, (double)mul32(q.E, add32(mul32(o.x, d.z), mul32(o.z, d.x))));
  b = add64(b, (double)mul32(q.F, add32(mul32(o.y, d.z), mul32(o.z, d.y))));
  b = add64(b, (double)mul32(q.G, d.x));
  b = add64(b, (double)mul32(q.H, d.y));
  b = add64(b, (double)mul32(q.I, d.z));
  float b_q = (float)b;
  if (a_q == 0.0f) {
    t = (float)__ddiv_rn(mul64(-1.0, (double)c_q), (double)b_q);
    return true;
  }
  float disc = (float)sub64(mul64((double)b_q, (double)b_q), mul64(mul64(4.0, (double)a_q), (double)c_q));
  if (disc < 0.0f) return false;
  double sq = sqrt((double)disc), den = mul64(2.0, (double)a_q), nb = -(double)b_q;
  t = (float)__ddiv_rn(sub64(nb, sq), den);
  if (t <= 0.0f) t = (float)__ddiv_rn(add64(nb, sq), den);
  return true;
}
