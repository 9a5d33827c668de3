// csrc/rt_qindex.cuh -- exact index over the ELLIPSOIDS among the quadrics (opt-in
// with rt_set_accel): which quadrics can be indexed, how far their bounding spheres must be inflated so
// that the index never drops a quadric the REFERENCE's arithmetic (C/raycast.c:614-656) would report,
// and the node test of the traversal.  Host + device code in one place: the device walks the hierarchy
// with qnode_prune(), the host builds it from qindex_classify(), and tests/test_qindex.py drives both
// through rt_dbg_qindex_* against the oracle's quadric intersector on adversarial rays.
//
// The reference evaluates, for origin P and direction d (floats, |d| = 1 up to 3 ulp),
//   a_q = fl(d'Md), b_q = fl(grad q(P).d), c_q = fl(q(P)),  hit <=> a_q == 0 or fl(b_q^2 - 4 a_q c_q) >= 0,
// the sums in double, but the products with d, e, f, g, h, i in float.  With u' = 2^-23 (twice the float
// unit: every constant below has that factor 2 in hand) and the exact values A, B, C of the same
// polynomials:   |a_q - A| <= ea = u'(A + 1.1 Sd),  |b_q - B| <= eb = u'(|B| + beta),  |c_q - C| <= ec = u'(|C| + gamma),
//   Sd = |d|+|e|+|f|,  S1 = |g|+|h|+|i|,  beta = S1 + 4.3 Sd om,  gamma = S1 om + 2 Sd om^2,  om = max |P component|.
// For an ellipsoid (M positive definite, eigenvalues lmin..lmax, centre c, q(c) = -K < 0) and s = |P - c|
// (so om <= |c|_inf + s, 0 < C <= lmax s^2, |B| <= 2 lmax s; the bounds become polynomials in s):
//  (1) MISS.  If the line stays outside {q <= delta}, then B^2 = 4A(C - qmin) with qmin > delta, and
//      b_q^2 < 4 a_q c_q follows from  delta >= u'[4C + 1.1 C Sd/lmin + beta sqrt(C/lmin) + gamma] + eb^2/(4 lmin)
//      <= a0 + a1 s + a2 s^2, so fl(disc) < 0: no hit.  {q <= delta} lies within sqrt((K + delta)/lmin) <=
//      r0 + rho s of the centre: r0 is the radius the hierarchy stores for the item, rho s is added by the
//      node test through its reach term (reach >= s).
//  (2) BEHIND.  If every point X of the bounding sphere has (P - X).d > m, both roots are below -m,
//      C = A t1 t2 >= A m |t1|, B > 2 A m > eb, and the reference returns a t <= 0 (no hit for a shadow
//      ray, rejected by the t > 0 rule for a nearest-hit search) provided 4 a_q c_q >= 2^-22 b_q^2, i.e.
//      m >= 2^-21 |t1| and m^2 >= 1.01 u' gamma / lmin:  m = behind_rel reach + behind_abs.
//  (3) BEYOND.  A reported t is >= t_entry - Dt, t_entry the entry into the bounding sphere,
//      Dt <= sqrt(Ed)/(2 a_q) + (half chord) ea/a_q + u'|t*|(2 + 1.1 Sd/lmin) + u' beta/(2 lmin),
//      Ed = |fl(disc) - D| <= u'(E2 s^2 + E0):  Dt <= t_rel reach + t_abs.
// A node of the hierarchy holds the bounding spheres of its members, and all margins grow with the node's
// radius and reach, so a node test that holds for the node sphere holds for every member.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

struct QIndexConsts {   // maxima over the ellipsoids of a node
  float miss_rel;             // bounding spheres grow by miss_rel * reach for the "line misses" test (float slack + rho)
  float behind_rel, behind_abs;
  float t_rel, t_abs;         // how far below the entry into its bounding sphere a reported t can fall
};

// Node (or item) test for the ray P + t d.  0 = enter, 1 = the line misses, 2 = wholly behind the origin,
// 3 = (nearest-hit search only) wholly beyond best_t.  nd = {centre, radius}.  *key (when the node is entered or
// beyond): the node holds no t below key / 1.001, so a traversal may drop it later once best_t * 1.001 < key.
__host__ __device__ inline int qnode_prune(const float4 nd, const float px, const float py, const float pz, const float dx,
                                           const float dy, const float dz, const bool nearest, const float best_t,
                                           const QIndexConsts k, float *key = nullptr) {
  const float wx = px - nd.x, wy = py - nd.y, wz = pz - nd.z;
  const float Tn = fmaf(wz, wz, fmaf(wy, wy, wx * wx));
  if (key) *key = -3.0e38f;
  if (!(Tn < 1e30f)) return 0;                                           // overflow or NaN origin: no pruning
  const float en = fmaf(dz, wz, fmaf(dy, wy, dx * wx));
  const float reach = sqrtf(Tn) + nd.w;                                  // >= |P - c| + r of every member
  const float rinf = fmaf(nd.w, 1.0001f, k.miss_rel * reach);
  if (!(Tn - en * en <= fmaf(rinf, rinf, 1e-5f * Tn))) return 1;        // 1e-5 Tn: float slack of this line, |d| != 1
  const float rbehind = fmaf(nd.w, 1.001f, k.behind_rel * reach) + k.behind_abs;
  if (en > rbehind) return 2;
  const float kk = (-en - rinf) - (k.t_rel * reach + k.t_abs);
  if (key) *key = kk;
  if (nearest && kk > best_t * 1.001f) return 3;
  return 0;
}

// ---- host side: which quadrics are ellipsoids the index can hold, and their constants -------------------
struct QIndexItem {
  bool ok;
  double c[3], K, lmin, lmax, a0, a1, a2;   // delta(s) <= a0 + a1 s + a2 s^2
  float centre[3], r_item;                  // bounding sphere of {q <= delta} without its rho s part, rounded outwards
  float miss_rel, behind_rel, behind_abs, t_rel, t_abs;   // this item's requirements (the scene uses the maxima)
};

// Eigenvalues of the symmetric 3x3 matrix {{a, d/2, e/2}, {d/2, b, f/2}, {e/2, f/2, c}} by cyclic Jacobi.
inline void qindex_eigen(const double a, const double b, const double c, const double d, const double e, const double f,
                         double ev[3]) {
  double m[3][3] = {{a, d / 2, e / 2}, {d / 2, b, f / 2}, {e / 2, f / 2, c}};
  for (int sweep = 0; sweep < 60; sweep++) {
    const double off = fabs(m[0][1]) + fabs(m[0][2]) + fabs(m[1][2]);
    if (off < 1e-300 || off <= 1e-18 * (fabs(m[0][0]) + fabs(m[1][1]) + fabs(m[2][2]))) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (m[p][q] == 0.0) continue;
        const double th = (m[q][q] - m[p][p]) / (2.0 * m[p][q]);
        const double t = (th >= 0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1.0));
        const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        for (int k = 0; k < 3; k++) { const double x = m[k][p], y = m[k][q]; m[k][p] = cs * x - sn * y; m[k][q] = sn * x + cs * y; }
        for (int k = 0; k < 3; k++) { const double x = m[p][k], y = m[q][k]; m[p][k] = cs * x - sn * y; m[q][k] = sn * x + cs * y; }
      }
  }
  ev[0] = m[0][0]; ev[1] = m[1][1]; ev[2] = m[2][2];
}

// q = {a, b, c, d, e, f, g, h, i, j} as the reference stores them (C/objects.h:31-40).  Returns out.ok = false
// for anything that is not a well-conditioned, finite, real ellipsoid.
inline void qindex_classify(const float *q, QIndexItem &out) {
  out.ok = false;
  for (int k = 0; k < 10; k++) if (!isfinite(q[k])) return;
  const double a = q[0], b = q[1], c = q[2], d = q[3], e = q[4], f = q[5], g = q[6], h = q[7], i = q[8], j = q[9];
  double ev[3];
  qindex_eigen(a, b, c, d, e, f, ev);
  const double lmin = fmin(ev[0], fmin(ev[1], ev[2])) * (1.0 - 1e-9), lmax = fmax(ev[0], fmax(ev[1], ev[2])) * (1.0 + 1e-9);
  if (!(lmin >= 1e-6) || !(lmax <= 1e6) || !(lmax / lmin <= 100.0)) return;   // also keeps every margin far above float underflow
  // centre: 2 M c = -(g, h, i)
  const double m00 = a, m01 = d / 2, m02 = e / 2, m11 = b, m12 = f / 2, m22 = c;
  const double det = m00 * (m11 * m22 - m12 * m12) - m01 * (m01 * m22 - m12 * m02) + m02 * (m01 * m12 - m11 * m02);
  if (!(det > 0.0)) return;
  const double r0 = -g / 2, r1 = -h / 2, r2 = -i / 2;
  const double cx = (r0 * (m11 * m22 - m12 * m12) - m01 * (r1 * m22 - m12 * r2) + m02 * (r1 * m12 - m11 * r2)) / det;
  const double cy = (m00 * (r1 * m22 - m12 * r2) - r0 * (m01 * m22 - m12 * m02) + m02 * (m01 * r2 - r1 * m02)) / det;
  const double cz = (m00 * (m11 * r2 - r1 * m12) - m01 * (m01 * r2 - r1 * m02) + r0 * (m01 * m12 - m11 * m02)) / det;
  // K = -q(c), evaluated directly with the cancellation it has, rounded up (a larger K only enlarges the bound)
  const double qc = a * cx * cx + b * cy * cy + c * cz * cz + d * cx * cy + e * cx * cz + f * cy * cz + g * cx + h * cy + i * cz + j;
  const double mag = fabs(a * cx * cx) + fabs(b * cy * cy) + fabs(c * cz * cz) + fabs(d * cx * cy) + fabs(e * cx * cz) +
                     fabs(f * cy * cz) + fabs(g * cx) + fabs(h * cy) + fabs(i * cz) + fabs(j);
  const double K = -qc + 1e-12 * mag;
  if (!(K > 0.0) || !isfinite(K) || !isfinite(cx) || !isfinite(cy) || !isfinite(cz)) return;
  if (!(K >= 1e-9 * mag)) return;                     // the surface must stand clear of the cancellation in q(c)
  const double up = ldexp(1.0, -23), SF = 1.5;
  const double Sd = fabs(d) + fabs(e) + fabs(f), S1 = fabs(g) + fabs(h) + fabs(i);
  const double cinf = fmax(fabs(cx), fmax(fabs(cy), fabs(cz)));
  const double kap = lmax / lmin, sk = sqrt(kap);
  if (!(up * (lmax + 1.1 * Sd) < 0.005 * lmin)) return;   // ea: a_q certainly positive, within 1 % of A
  // om <= cinf + s:  beta <= b0 + 4.3 Sd s,  gamma <= g0 + g1 s + g2 s^2
  const double b0 = S1 + 4.3 * Sd * cinf;
  const double g0 = S1 * cinf + 2.0 * Sd * cinf * cinf, g1 = S1 + 4.0 * Sd * cinf, g2 = 2.0 * Sd;
  // (1) delta(s) <= a0 + a1 s + a2 s^2
  const double a2 = SF * (up * ((4.0 + 1.1 * Sd / lmin) * lmax + 4.3 * Sd * sk + g2) + up * up * (2.0 * lmax * lmax + 18.5 * Sd * Sd) / lmin);
  const double a1 = SF * up * (b0 * sk + g1);
  const double a0 = SF * (up * g0 + up * up * b0 * b0 / lmin);
  const double eps = ldexp(1.0, -8);
  const double r_0 = sqrt((K + a0) / lmin) + a1 / (2.0 * lmin * eps);       // sqrt(a1 s / lmin) <= a1/(2 lmin eps) + eps s/2
  const double rho = sqrt(a2 / lmin) + eps / 2.0;
  // (2) behind:  m >= sqrt(2 u' gamma/lmin) + 4 u' beta/lmin + 2^-21 reach
  const double behind_abs = SF * (sqrt(2.0 * up * g0 / lmin) + up * g1 / (lmin * eps) + 4.0 * up * b0 / lmin);
  const double behind_rel = ldexp(1.0, -6) + SF * (sqrt(2.0 * up * g2 / lmin) + eps / 2.0 + 17.2 * up * Sd / lmin);
  // (3) beyond
  const double E2 = lmax * (28.0 * lmax + 20.4 * Sd) + 74.0 * Sd * Sd;
  const double E0 = 4.0 * b0 * b0 + 2.0 * S1 * S1 + 12.0 * lmax * K + 4.4 * K * Sd + 4.0 * lmax * (S1 * cinf + 4.0 * Sd * cinf * cinf);
  const double t_rel = SF * (sqrt(up * E2) / (1.98 * lmin) + up * (2.0 + 1.1 * Sd / lmin) + 2.15 * up * Sd / lmin);
  const double t_abs = SF * (sqrt(up * E0) / (1.98 * lmin) + sqrt(K / lmin) * up * (kap + 1.1 * Sd / lmin) * 1.05 + up * b0 / (2.0 * lmin));
  if (!isfinite(r_0) || !isfinite(t_abs) || !isfinite(behind_abs) || !(t_rel < 0.125) || !(rho < 0.0625) || !(behind_rel < 0.125)) return;
  out.c[0] = cx; out.c[1] = cy; out.c[2] = cz; out.K = K; out.lmin = lmin; out.lmax = lmax;
  out.a0 = a0; out.a1 = a1; out.a2 = a2;
  out.centre[0] = (float)cx; out.centre[1] = (float)cy; out.centre[2] = (float)cz;
  const double shift = sqrt((cx - out.centre[0]) * (cx - out.centre[0]) + (cy - out.centre[1]) * (cy - out.centre[1]) +
                            (cz - out.centre[2]) * (cz - out.centre[2]));
  out.r_item = (float)((r_0 + shift) * (1.0 + 1e-6)) + 1e-30f;
  out.miss_rel = (float)((0.0029296875 + rho) * (1.0 + 1e-6));             // 2^-9 + 2^-10: float slack of the node test, generous
  out.behind_rel = (float)(behind_rel * (1.0 + 1e-6));
  out.behind_abs = (float)(behind_abs * (1.0 + 1e-6));
  out.t_rel = (float)(t_rel * (1.0 + 1e-6));
  out.t_abs = (float)(t_abs * (1.0 + 1e-6));
  if (!isfinite(out.r_item)) return;
  out.ok = true;
}
