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

// ---- conservative FP32 reject filter for quadrics (DESIGN.md "quadric filter") -------
// The reference (C/raycast.c:614-656) misses a quadric when a_q != 0 and
// disc = b_q^2 - 4 a_q c_q < 0.  With FP32/FMA estimates a', b', c' of the three
// coefficients and bounds on what separates them from the reference's floats
//   |a_q - a'| <= ea = 2^-19 S2,   |b_q - b'| <= eb = 2^-19 (2 S2 om + S1),
//   |c_q - c'| <= ec = 2^-19 (S2 om^2 + S1 om + |J|)
// (S2 = |A|+..+|F|, S1 = |G|+|H|+|I|, om = max |origin component|, |d| = 1; the true
// separations are below 12 u times the same sums), a ray certainly misses when
//   (|b'| + eb)^2  <  4 (|c'| - ec)(1 - 2^-18) (sigma a' - ea),   sigma = sign(c'),
// because then sign(a_q) = sign(c_q), a_q != 0 and b_q^2 < 4 a_q c_q with a relative gap of
// 2^-18.  Everything that depends only on the origin is folded into scaled coefficients,
// so one ray costs 9 + 3 + 2 FMA-pipe instructions:  v = beta^2 - m,  v < 0  =>  miss.
#define RT_QUAD_EPS 1.9073486328125e-06f   /* 2^-19 */
struct QuadFilter {
  float As, Bs, Cs, Ds, Es, Fs, m0;   // K4 sigma * (A..F), -K4 ea
  float gx, gy, gz, eb;               // gradient estimate at the origin, bound on b
};
// q0 = {A,B,C,D}, q1 = {E,F,G,H}, q2 = {I,J,S2,S1}
__host__ __device__ inline QuadFilter quad_filter_setup(float4 q0, float4 q1, float4 q2, float ox, float oy, float oz) {
  const float A = q0.x, B = q0.y, C = q0.z, D = q0.w, E = q1.x, F = q1.y, G = q1.z, H = q1.w, I = q2.x, J = q2.y;
  const float S2 = q2.z, S1 = q2.w;
  const float om = fmaxf(fabsf(ox), fmaxf(fabsf(oy), fabsf(oz)));
  QuadFilter f;
  f.gx = fmaf(2.0f * A, ox, fmaf(D, oy, fmaf(E, oz, G)));
  f.gy = fmaf(D, ox, fmaf(2.0f * B, oy, fmaf(F, oz, H)));
  f.gz = fmaf(E, ox, fmaf(F, oy, fmaf(2.0f * C, oz, I)));
  const float u1 = fmaf(A, ox, fmaf(D, oy, fmaf(E, oz, G))), u2 = fmaf(B, oy, fmaf(F, oz, H)), u3 = fmaf(C, oz, I);
  const float cp = fmaf(ox, u1, fmaf(oy, u2, fmaf(oz, u3, J)));
  const float cabs = fmaf(S2, om * om, fmaf(S1, om, fabsf(J)));
  f.eb = RT_QUAD_EPS * fmaf(2.0f * S2, om, S1);
  const float C1 = fabsf(cp) - RT_QUAD_EPS * cabs;
  // K4 = 0 switches the filter off for this quadric and origin (everything survives):
  // c too close to 0 to know its sign, or magnitudes where the scaled products could overflow.
  float K4 = 0.0f;
  if (C1 > 0.0f && C1 < 1e15f && S2 < 1e15f) K4 = copysignf(4.0f * C1 * (1.0f - 3.814697265625e-06f), cp);
  f.As = K4 * A; f.Bs = K4 * B; f.Cs = K4 * C; f.Ds = K4 * D; f.Es = K4 * E; f.Fs = K4 * F;
  f.m0 = -fabsf(K4) * (RT_QUAD_EPS * 1.0001f * S2);
  return f;
}
// v >= 0 (or NaN-free inputs that cannot be rejected) => the ray survives the filter.
RT_DEV float quad_filter_v(const QuadFilter &f, v3 d) {
  const float u1 = __fmaf_rn(f.Es, d.z, __fmaf_rn(f.Ds, d.y, f.As * d.x));
  const float u2 = __fmaf_rn(f.Fs, d.z, f.Bs * d.y);
  const float u3 = f.Cs * d.z;
  const float m = __fmaf_rn(d.x, u1, __fmaf_rn(d.y, u2, __fmaf_rn(u3, d.z, f.m0)));
  const float b = __fmaf_rn(f.gz, d.z, __fmaf_rn(f.gy, d.y, f.gx * d.x));
  const float beta = fabsf(b) + f.eb;
  return __fmaf_rn(beta, beta, -m);
}
