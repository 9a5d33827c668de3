/*
 * oracle/rt_oracle.c -- TEST INFRASTRUCTURE, never shipped, never measured as product.
 *
 * CPU restatement of the reference's per-pixel render loop (the PATCHED semantics of
 * P1 num_lights zeroed, P2 next_nearest_object reset), written
 * from the behaviour of /root/reference/C/raycast.c and C/v3math.c, over the flat
 * rt_scene of include/rt_abi.h.  Each function cites the reference lines it follows.
 * Typing is literal: an expression the reference evaluates in double is evaluated in
 * double here, float stays float; build with -ffp-contract=off and never -ffast-math.
 * Two deliberate, documented deviations from libm (shared bit-for-bit with the CUDA path):
 *   - pow(x, 0.5)  -> sqrt(x)                 (C/raycast.c:593,596,649,651)
 *   - pow(angle,20) and pow(alpha, small integer a0) -> fixed multiplication chains
 *     (C/raycast.c:695,755-757); non-integer a0 still calls pow().
 * Parity is PINNED against the reference itself: tests/test_oracle_vs_reference.py
 * compares this file with oracle/_ref/libref.so (the patched reference compiled from
 * /root/reference) and tests/golden/ holds images generated from that library.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include "rt_abi.h"

static int g_max_level = 7; /* MAX_RECURSION, C/raycast.c:14; the reference's CUDA/OptiX variants run the same loop
                               with MAX_ITER 50 (CUDA/raycast.cu:13): orc_set_max_level(50) restates that dormant mode */
#define MAX_LEVEL g_max_level
void orc_set_max_level(int n) { g_max_level = n < 1 ? 1 : n; }

typedef struct orc_counters {
  uint64_t rays_primary, rays_reflection, rays_shadow, hit_points;
  uint64_t tests_sphere, tests_plane, tests_quadric; /* reference-loop tests (no early exit) */
  uint64_t shadow_occluded;
} orc_counters;

/* ---- v3math (C/v3math.c) ------------------------------------------------ */
static inline float dot3(const float *a, const float *b) { /* :69-71 */
  return (a[0] * b[0]) + (a[1] * b[1]) + (a[2] * b[2]);
}
static inline float len3(const float *a) { /* :169-172: double squares and sums, double sqrt */
  double s = (double)a[0] * (double)a[0] + (double)a[1] * (double)a[1] + (double)a[2] * (double)a[2];
  return (float)sqrt(s);
}
static inline void norm3(float *a) { /* :180-192: no-op on zero length, three divisions */
  float l = len3(a);
  if (l == 0.0f) return;
  a[0] = a[0] / l; a[1] = a[1] / l; a[2] = a[2] / l;
}
static inline void reflect3(float *dst, const float *v, const float *n) { /* :144-160 */
  float s = 2 * dot3(v, n);
  float p0 = n[0] * s, p1 = n[1] * s, p2 = n[2] * s;
  dst[0] = v[0] - p0; dst[1] = v[1] - p1; dst[2] = v[2] - p2;
}

/* x^20 as a fixed chain of double multiplications (shared with the CUDA path). */
static inline double pow20(double x) {
  double x2 = x * x, x4 = x2 * x2, x5 = x4 * x, x10 = x5 * x5;
  return x10 * x10;
}
/* alpha^a0: integer 0..64 exponents by binary exponentiation, else libm pow. */
static inline double pow_a0(double x, float a0) {
  if (a0 >= 0.0f && a0 <= 64.0f && (float)(int)a0 == a0) {
    int e = (int)a0; double r = 1.0, b = x;
    while (e) { if (e & 1) r = r * b; b = b * b; e >>= 1; }
    return r;
  }
  return pow(x, (double)a0);
}

/* ---- intersections ------------------------------------------------------ */
/* C/raycast.c:576-600 */
static int isect_sphere(const float *o, const float *d, const rt_shape *s, float *t) {
  float tmp[3] = {o[0] - s->position[0], o[1] - s->position[1], o[2] - s->position[2]};
  float r = s->params[0];
  float a = (float)((double)d[0] * (double)d[0] + (double)d[1] * (double)d[1] + (double)d[2] * (double)d[2]);
  float b = 2 * dot3(d, tmp);
  float c = (float)((double)dot3(tmp, tmp) - (double)r * (double)r);
  float disc = (float)((double)b * (double)b - (double)(4 * a * c));
  if (disc < 0) return 0;
  *t = (float)((-(double)b - sqrt((double)disc)) / (2.0 * (double)a));
  if (*t < 0) *t = (float)((-(double)b + sqrt((double)disc)) / (2.0 * (double)a));
  return 1;
}
/* C/raycast.c:545-562 */
static int isect_plane(const float *o, const float *d, const rt_shape *s, float *t) {
  float sub[3] = {o[0] - s->position[0], o[1] - s->position[1], o[2] - s->position[2]};
  float num = dot3(sub, s->params);
  float den = dot3(d, s->params);
  if (den == 0) return 0;
  float tt = -1 * num / den;
  if (tt < 0) return 0;
  *t = tt;
  return 1;
}
/* C/raycast.c:614-656. Terms without a double factor are float products first. */
static int isect_quadric(const float *o, const float *d, const rt_shape *s, float *t) {
  const float *q = s->params;
  const float A = q[0], B = q[1], C = q[2], D = q[3], E = q[4], F = q[5], G = q[6], H = q[7], I = q[8], J = q[9];
  double d0s = (double)d[0] * (double)d[0], d1s = (double)d[1] * (double)d[1], d2s = (double)d[2] * (double)d[2];
  double o0s = (double)o[0] * (double)o[0], o1s = (double)o[1] * (double)o[1], o2s = (double)o[2] * (double)o[2];
  float a_q = (float)((double)A * d0s + (double)B * d1s + (double)C * d2s
                      + (double)(D * d[0] * d[1]) + (double)(E * d[0] * d[2]) + (double)(F * d[1] * d[2]));
  float b_q = (float)(2.0 * (double)A * (double)o[0] * (double)d[0]
                      + 2.0 * (double)B * (double)o[1] * (double)d[1]
                      + 2.0 * (double)C * (double)o[2] * (double)d[2]
                      + (double)(D * (o[0] * d[1] + o[1] * d[0]))
                      + (double)(E * (o[0] * d[2] + o[2] * d[0]))
                      + (double)(F * (o[1] * d[2] + o[2] * d[1]))
                      + (double)(G * d[0]) + (double)(H * d[1]) + (double)(I * d[2]));
  float c_q = (float)((double)A * o0s + (double)B * o1s + (double)C * o2s
                      + (double)(D * o[0] * o[1]) + (double)(E * o[0] * o[2]) + (double)(F * o[1] * o[2])
                      + (double)(G * o[0]) + (double)(H * o[1]) + (double)(I * o[2]) + (double)J);
  if (a_q == 0.0) {
    *t = (float)(-1.0 * (double)c_q / (double)b_q);
  } else {
    float disc = (float)((double)b_q * (double)b_q - 4.0 * (double)a_q * (double)c_q);
    if (disc < 0.0) return 0;
    *t = (float)((-(double)b_q - sqrt((double)disc)) / (2.0 * (double)a_q));
    if (*t <= 0) *t = (float)((-(double)b_q + sqrt((double)disc)) / (2.0 * (double)a_q));
  }
  return 1;
}

/* C/raycast.c:441-531. Returns the index (-1 = none); P/N only when !shadow. */
static int nearest(const rt_scene *sc, const float *o, const float *d, float *P, float *N,
                   int skip, int shadow, orc_counters *ct) {
  float best = INFINITY, t;
  int idx = -1;
  for (int i = 0; i < sc->num_shapes; i++) {
    if (i == skip) continue;
    const rt_shape *s = &sc->shapes[i];
    if (s->type == RT_SPHERE) {
      if (ct) ct->tests_sphere++;
      if (!isect_sphere(o, d, s, &t)) continue;
      if (!(best > t && t > 0)) continue;
      best = t; idx = i;
      if (!shadow) {
        for (int k = 0; k < 3; k++) P[k] = o[k] + d[k] * best;
        float inv = (float)(1.0 / (double)s->params[0]);
        for (int k = 0; k < 3; k++) N[k] = (P[k] - s->position[k]) * inv;
        norm3(N);
      }
    } else if (s->type == RT_PLANE) {
      if (ct) ct->tests_plane++;
      if (!isect_plane(o, d, s, &t)) continue;
      if (!(best > t && t > 0)) continue;
      best = t; idx = i;
      if (!shadow) {
        for (int k = 0; k < 3; k++) { P[k] = o[k] + d[k] * best; N[k] = s->params[k]; }
      }
    } else {
      if (ct) ct->tests_quadric++;
      if (!isect_quadric(o, d, s, &t)) continue;
      if (skip != -1 && (o[2] + t * d[2]) < o[2]) continue; /* :492-494 */
      if (!(best > t && t > 0)) continue;
      best = t; idx = i;
      if (!shadow) {
        const float *q = s->params;
        for (int k = 0; k < 3; k++) P[k] = o[k] + d[k] * best;
        N[0] = (float)(2.0 * (double)q[0] * (double)P[0] + (double)(q[3] * P[1]) + (double)(q[4] * P[2]) + (double)q[6]);
        N[1] = (float)(2.0 * (double)q[1] * (double)P[1] + (double)(q[3] * P[0]) + (double)(q[5] * P[2]) + (double)q[7]);
        N[2] = (float)(2.0 * (double)q[2] * (double)P[2] + (double)(q[4] * P[0]) + (double)(q[5] * P[1]) + (double)q[8]);
        norm3(N);
        if (dot3(N, d) > 0) { N[0] = N[0] * -1.0f; N[1] = N[1] * -1.0f; N[2] = N[2] * -1.0f; }
      }
    }
  }
  return idx;
}

/* ---- shading ------------------------------------------------------------ */
/* C/raycast.c:381-421 with the helpers :666-758 inlined in their own typing. */
static void shade(const rt_scene *sc, float *out, int oi, const float *P, const float *N,
                  const float *V, orc_counters *ct) {
  const rt_shape *ob = &sc->shapes[oi];
  float opacity = (float)(1.0 - (double)ob->reflectivity - (double)ob->refractivity); /* :383 */
  out[0] = out[1] = out[2] = 0;
  if (ct) ct->hit_points++;
  if (!(opacity > 0.0)) return;
  for (int l = 0; l < sc->num_lights; l++) {
    const rt_light *lt = &sc->lights[l];
    float Ld[3] = {lt->position[0] - P[0], lt->position[1] - P[1], lt->position[2] - P[2]}; /* :396 */
    float dist = len3(Ld);
    norm3(Ld);
    if (ct) ct->rays_shadow++;
    if (nearest(sc, P, Ld, NULL, NULL, oi, 1, ct) != -1) { if (ct) ct->shadow_occluded++; continue; } /* :401-405 */
    /* radial_attenuation :666-669 */
    float rad = (float)(1.0 / ((double)(lt->radial[0] + lt->radial[1] * dist)
                               + (double)lt->radial[2] * ((double)dist * (double)dist)));
    /* angular_attenuation :679-696 */
    float ang;
    if (lt->type != RT_SPOTLIGHT) ang = 1.0f;
    else {
      float vo[3] = {P[0] - lt->position[0], P[1] - lt->position[1], P[2] - lt->position[2]};
      norm3(vo);
      float alpha = dot3(vo, lt->direction);
      if (alpha < lt->cos_theta) ang = 0.0f;
      else ang = (float)pow_a0((double)alpha, lt->a0);
    }
    /* diffuse_light :708-720, specular_light :733-758 */
    float diff[3] = {0, 0, 0}, spec[3] = {0, 0, 0};
    float theta = dot3(N, Ld);
    if (!(theta <= 0.0)) {
      for (int k = 0; k < 3; k++) diff[k] = ob->diffuse[k] * lt->color[k] * theta;
      float R[3];
      reflect3(R, Ld, N);
      float nV[3] = {V[0] * -1.0f, V[1] * -1.0f, V[2] * -1.0f};
      double angle = dot3(nV, R);
      if (!(angle > 0)) {
        double p = pow20(angle);
        for (int k = 0; k < 3; k++) spec[k] = (float)((double)(ob->specular[k] * lt->color[k]) * p);
      }
    }
    for (int k = 0; k < 3; k++) out[k] += (diff[k] + spec[k]) * rad * ang; /* :414-416 */
  }
  for (int k = 0; k < 3; k++) out[k] = out[k] * opacity; /* :419 */
}

/* C/raycast.c:315-379 (patched: a missed reflection ray ends the chain). */
static void trace(const rt_scene *sc, const float *o, const float *d, float *out, orc_counters *ct) {
  float P0[3], N0[3], c[3];
  out[0] = out[1] = out[2] = 0;
  if (ct) ct->rays_primary++;
  int i0 = nearest(sc, o, d, P0, N0, -1, 0, ct);
  if (i0 < 0) return;
  float cd[3] = {d[0], d[1], d[2]}, co[3] = {P0[0], P0[1], P0[2]}, cn[3] = {N0[0], N0[1], N0[2]};
  float P[3], N[3];
  int skip = i0, cur = i0, level = 1;
  float tr = sc->shapes[i0].reflectivity;
  while (level < MAX_LEVEL) {
    if (!(sc->shapes[cur].reflectivity > 0)) break;
    float rv[3];
    reflect3(rv, cd, cn);
    norm3(rv);
    cd[0] = rv[0]; cd[1] = rv[1]; cd[2] = rv[2];
    if (ct) ct->rays_reflection++;
    int hi = nearest(sc, co, cd, P, N, skip, 0, ct);
    if (hi < 0) break;
    shade(sc, c, hi, P, N, cd, ct);
    for (int k = 0; k < 3; k++) c[k] = c[k] * tr;
    tr *= sc->shapes[hi].reflectivity;
    for (int k = 0; k < 3; k++) { out[k] = out[k] + c[k]; co[k] = P[k]; cn[k] = N[k]; }
    skip = hi; cur = hi; level++;
  }
  shade(sc, c, i0, P0, N0, d, ct);
  for (int k = 0; k < 3; k++) out[k] = out[k] + c[k];
}

/* The dormant recursive tracer shoot() (C/raycast.c:152-296), orc_set_mode(1).  What it RETURNS: its refraction
 * subtree (:186-243) is traced but the result is never added (the two lines that would are commented out, :293,295),
 * so the value is the nested reflection sum
 *     out_k = colour_k * opacity_k  +  out_(k+1) * reflectivity_k      (:291-294; out beyond the chain's end = black),
 * with the same hits, shadow rays and Phong terms as the live iterative_shoot -- only the order of the float
 * additions and multiplications differs (nested from the deepest hit outwards instead of a running product).
 * recursion_level starts at MAX_LEVEL (7) and a call at level 0 returns black (:156-159): primary + 6 bounces, as
 * the live path.  The discarded refraction rays are not traced here (dead code; in the reference they read
 * uninitialised memory when the inner ray misses, :214-221). */
static void trace_shoot(const rt_scene *sc, const float *o, const float *d, float *out, int level, int skip, orc_counters *ct) {
  out[0] = out[1] = out[2] = 0;
  if (level == 0) return;
  float P[3], N[3];
  if (ct) { if (skip < 0) ct->rays_primary++; else ct->rays_reflection++; }
  const int hi = nearest(sc, o, d, P, N, skip, 0, ct);
  if (hi < 0) return;
  const rt_shape *ob = &sc->shapes[hi];
  float rc[3] = {0, 0, 0};
  if (ob->reflectivity > 0) {
    float rv[3];
    reflect3(rv, d, N);
    norm3(rv);
    if (level - 1 > 0) trace_shoot(sc, P, rv, rc, level - 1, hi, ct);
  }
  float c[3];
  shade(sc, c, hi, P, N, d, ct);                        /* colour * opacity, :244-290 = calc_color */
  for (int k = 0; k < 3; k++) out[k] = c[k] + rc[k] * ob->reflectivity;   /* :291-294 */
}
static int g_mode = 0;   /* 0: iterative_shoot (the live path), 1: shoot */
void orc_set_mode(int m) { g_mode = m ? 1 : 0; }

static inline float clampf(float v, float lo, float hi) { /* ppm_clamp, C/ppm.c:350-359 */
  if (v > hi) v = hi;
  if (v < lo) v = lo;
  return v;
}
static inline uint8_t quant(float v) {
  float c = clampf(v, 0, 255);
  return (c == c) ? (uint8_t)c : 0; /* NaN: x86 cvttss2si low byte and __float2uint_rz both give 0 */
}

/* Rows [y0,y1) of C/raycast.c:109-129. */
void orc_render_rows(const rt_scene *sc, int W, int H, int y0, int y1, uint8_t *out, orc_counters *ct) {
  float o[3] = {0, 0, 0}, d[3], color[3];
  float ph = sc->camera_height / H, pw = sc->camera_width / W;
  uint8_t *px = out;
  for (int y = y0; y < y1; y++) {
    for (int x = 0; x < W; x++) {
      d[0] = (float)(0 - ((double)sc->camera_width / 2.0) + (double)pw * (x + 0.5));
      d[1] = (float)(0 + ((double)sc->camera_height / 2.0) - (double)ph * (y + 0.5));
      d[2] = -1;
      norm3(d);
      if (g_mode) trace_shoot(sc, o, d, color, MAX_LEVEL, -1, ct);
      else trace(sc, o, d, color, ct);
      for (int k = 0; k < 3; k++) px[k] = quant(color[k] * 255);
      px += 3;
    }
  }
}

typedef struct { const rt_scene *sc; int W, H; const int *rows; int n; uint8_t *out; int tid, nthr; orc_counters ct; } orc_job;
static void *orc_worker(void *a) {
  orc_job *j = (orc_job *)a;
  for (int i = j->tid; i < j->n; i += j->nthr)
    orc_render_rows(j->sc, j->W, j->H, j->rows[i], j->rows[i] + 1, j->out + (size_t)i * j->W * 3, &j->ct);
  return NULL;
}
/* rows[i] -> out + i*W*3, over nthreads host threads; counters summed into *ct (may be NULL). */
void orc_render_row_list(const rt_scene *sc, int W, int H, const int *rows, int n, uint8_t *out,
                         int nthreads, orc_counters *ct) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  pthread_t th[256];
  orc_job *jb = (orc_job *)calloc((size_t)nthreads, sizeof(orc_job));
  for (int t = 0; t < nthreads; t++) {
    jb[t].sc = sc; jb[t].W = W; jb[t].H = H; jb[t].rows = rows; jb[t].n = n; jb[t].out = out;
    jb[t].tid = t; jb[t].nthr = nthreads;
    pthread_create(&th[t], NULL, orc_worker, &jb[t]);
  }
  for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
  if (ct) {
    memset(ct, 0, sizeof *ct);
    for (int t = 0; t < nthreads; t++) {
      uint64_t *dst = (uint64_t *)ct, *src = (uint64_t *)&jb[t].ct;
      for (size_t k = 0; k < sizeof(orc_counters) / sizeof(uint64_t); k++) dst[k] += src[k];
    }
  }
  free(jb);
}

/* Known-answer entry points mirroring ref_isect_* of ref_harness.c. */
int orc_isect_sphere(const float *o, const float *d, const float *centre, float r, float *t) {
  rt_shape s; memset(&s, 0, sizeof s); s.type = RT_SPHERE; s.params[0] = r;
  for (int k = 0; k < 3; k++) s.position[k] = centre[k];
  return isect_sphere(o, d, &s, t);
}
int orc_isect_plane(const float *o, const float *d, const float *pos, const float *n, float *t) {
  rt_shape s; memset(&s, 0, sizeof s); s.type = RT_PLANE;
  for (int k = 0; k < 3; k++) { s.position[k] = pos[k]; s.params[k] = n[k]; }
  return isect_plane(o, d, &s, t);
}
int orc_isect_quadric(const float *o, const float *d, const float *q10, float *t) {
  rt_shape s; memset(&s, 0, sizeof s); s.type = RT_QUADRIC;
  for (int k = 0; k < 10; k++) s.params[k] = q10[k];
  return isect_quadric(o, d, &s, t);
}

/* n rays against one quadric (tests/test_qindex.py): hit[r] = the intersector's return value, t[r] its distance */
void orc_isect_quadric_batch(const float *o3, const float *d3, const float *q10, int n, int *hit, float *t) {
  rt_shape s; memset(&s, 0, sizeof s); s.type = RT_QUADRIC;
  for (int k = 0; k < 10; k++) s.params[k] = q10[k];
  for (int r = 0; r < n; r++) { t[r] = 0.0f; hit[r] = isect_quadric(o3 + 3 * r, d3 + 3 * r, &s, &t[r]); }
}

/* ---- P3 writer restatement (C/ppm.c:168-184): "P3\n", "%d %d \n%u\n", then per pixel
 * "%d\n%d\n%d\n" of the three channel bytes, rows top to bottom.  Writes into out (capacity
 * cap) and returns the length, or 0 when cap is too small.  size_t indexing (the reference's
 * int arithmetic is the same below 2^31 values). */
#include <stdio.h>
size_t orc_p3_format(const uint8_t *pixmap, int width, int height, unsigned max_color, char *out, size_t cap) {
  size_t n = 0;
  if (cap < 64) return 0;
  n += (size_t)sprintf(out + n, "P3\n");
  n += (size_t)sprintf(out + n, "%d %d \n%u\n", width, height, (unsigned)(uint8_t)max_color);
  for (int y = 0; y < height; y++) {
    for (int x = 0; x < width; x++) {
      const uint8_t *pixel = &pixmap[3 * ((size_t)y * (size_t)width + (size_t)x)];
      if (n + 16 > cap) return 0;
      n += (size_t)sprintf(out + n, "%d\n%d\n%d\n", pixel[0], pixel[1], pixel[2]);
    }
  }
  return n;
}
