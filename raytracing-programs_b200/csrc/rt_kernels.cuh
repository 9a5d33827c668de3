// csrc/rt_kernels.cuh -- the render-loop kernels (sm_100a).  See DESIGN.md for the
// wavefront structure:  k_primary -> k_shade<LB> x 7 levels -> k_finalize.
//   k_primary  : C/raycast.c:113-121 + the first find_nearest_object (:325)
//   k_shade    : calc_color (:381-421) for one hit point, i.e. its L shadow rays, plus
//                the reflection ray leaving the same point (:352-358): all 1+L rays
//                share origin and skip index, so the origin-only part of every
//                intersection test is computed once per primitive and reused.
//   k_finalize : the acc + primary sum (:377-378), *255, clamp, truncate (:122-126),
//                stored 4 pixels = 12 bytes at a time.
#pragma once
#include "rt_scene.cuh"

#ifndef RT_SHADE_MINB
#define RT_SHADE_MINB 4
#endif
#ifndef RT_SHADE_MINB_NARROW
#define RT_SHADE_MINB_NARROW RT_SHADE_MINB   /* CTAs per SM of the instantiations with at most 8 lights per pass (9 ray slots) */
#endif
#ifndef RT_VSTEP
#define RT_VSTEP 2       /* the bundle scans are compiled for NR, NR - RT_VSTEP, ... ray slots.  The instruction caches
                            decide (L0 ~6 KB per scheduler, L1.5 32 KB per SM): with 4-sphere trips and the resolve
                            stage inlined per variant, step 2 was 37 % SLOWER than step 4 (13.5 k instructions);
                            with 2-sphere trips and one shared resolve function it is 5 % faster (4.8 k) */
#endif
#ifndef RT_SCAN4_MAX
#define RT_SCAN4_MAX 0   /* scan variants up to this many slots take 4 spheres per loop trip, wider ones 2 */
#endif
#ifndef RT_TRIP1_MIN
#define RT_TRIP1_MIN 13  /* constant-bank scan: variants with >= this many slots take 1 sphere per loop trip, */
#endif
#ifndef RT_PREFETCH2
#define RT_PREFETCH2 0
#endif
#ifndef RT_TRIP2_MIN
#define RT_TRIP2_MIN 5   /* >= this many 2, narrower ones 4 */
#endif
#define RT_CHUNK 128     /* spheres per scan chunk = capacity of the per-thread survivor queue */
#define RT_INF __int_as_float(0x7f800000)
#define RT_NAN __int_as_float(0x7fc00000)
#define RT_FILTER_EPS 3.814697265625e-06f   /* 2^-18, margin of the sphere reject filter */
#define RT_P2_EPS 1.9073777e-06f            /* 2^-19 (1 + 2^-16): weight of |P|^2 in the margin of the constant-bank scan */

struct TileGeom {
  int width, height;        // full image
  int rank, nranks, band_h; // row interleave: band b -> rank b % nranks
  long long first_px;       // index of the tile's first pixel in the rank buffer (local_row0 * width)
  int n_px;                 // pixels in this tile
};

RT_DEV int local_to_image_row(int lr, int rank, int nranks, int band_h) {
  int band = lr / band_h;
  return (band * nranks + rank) * band_h + (lr - band * band_h);
}

RT_DEV bool better_hit(float t, int idx, float best_t, int best_idx) {
  // nearest_distance > distance && distance > 0 (C/raycast.c:456) with the reference's
  // lowest-index-wins tie rule made order independent: lexicographic (t, index).
  return (t > 0.0f) && (best_t > t || (best_t == t && idx < best_idx && t < RT_INF));
}

RT_DEV int warp_agg_add(int *counter) {
  unsigned m = __activemask();
  int lane = threadIdx.x & 31;
  int leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(counter, __popc(m));
  base = __shfl_sync(m, base, leader);
  return base + __popc(m & ((1u << lane) - 1u));
}

// Work class of a new hit point: how many ray slots its shading will need (estimate).
// Same rule as k_shade's ray setup, but on the un-normalised light vector and without
// caring about rounding: a wrong guess only costs balance, never correctness.
RT_DEV int hit_class(const DevScene &sc, v3 P, v3 N, int obj, bool can_reflect, int nr_max) {
  if (nr_max < 3 + RT_VSTEP) return 0;    // one scan variant only: nothing to sort
  const float refl = sc.mat[2 * obj].w, opacity = sc.mat[2 * obj + 1].w;
  int cnt = (can_reflect && refl > 0.0f) ? 1 : 0;
  if (opacity > 0.0f) {
    const int nl = min(sc.n_lights, nr_max - 1);
    for (int l = 0; l < nl; l++) {
      const float4 lp = sc.lights[4 * l];
      const float s = __fmaf_rn(N.z, lp.z - P.z, __fmaf_rn(N.y, lp.y - P.y, N.x * (lp.x - P.x)));
      cnt += (sc.lights[4 * l + 3].w == 0.0f || !(s <= 0.0f)) ? 1 : 0;
    }
  }
  // one class per scan variant of k_shade (nr_max, nr_max - RT_VSTEP, ... >= 3 slots), the narrowest = class 0
  const int vmin = nr_max - RT_VSTEP * ((nr_max - 3) / RT_VSTEP);
  int v = nr_max - RT_VSTEP * ((nr_max - cnt) / RT_VSTEP);
  if (v < vmin) v = vmin;
  const int c = (v - vmin) / RT_VSTEP;
  return c < RT_NCLS ? c : RT_NCLS - 1;
}

// Appends one entry per lane that has `on` set.  Called by ALL 32 lanes in converged code (full-mask collectives, no
// __activemask()): collectives on a partial mask inside a divergent branch left ptxas unable to prove the warp
// converged anywhere in the kernel, and without that proof it keeps the scan off the uniform datapath.
RT_DEV void queue_push(const HitQueue &q, const bool on, float4 a, float4 b, float4 c, int cls) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const unsigned act = __ballot_sync(0xffffffffu, on);
  int base = 0;
  if (lane == 0 && act) base = atomicAdd(&q.ctl[0], __popc(act));
  base = __shfl_sync(0xffffffffu, base, 0);
  const int slot = base + __popc(act & lt);
  if (on) { q.r0[slot] = a; q.r1[slot] = b; q.r2[slot] = c; }
  // class list position, aggregated over the lanes that file under the same class so that
  // neighbouring pixels stay neighbours in the list (their shading shares survivors and cache lines)
  #pragma unroll
  for (int k = 0; k < RT_NCLS; k++) {
    const unsigned peers = __ballot_sync(0xffffffffu, on && cls == k);
    int cb = 0;
    if (lane == 0 && peers) cb = atomicAdd(&q.ctl[1 + k], __popc(peers));
    cb = __shfl_sync(0xffffffffu, cb, 0);
    if (on && cls == k) q.cls[(size_t)k * q.cap + cb + __popc(peers & lt)] = slot;
  }
}

RT_DEV quad10 load_quad(const float4 *q, int i) {
  float4 a = q[3 * i], b = q[3 * i + 1], c = q[3 * i + 2];
  quad10 r;
  r.A = a.x; r.B = a.y; r.C = a.z; r.D = a.w; r.E = b.x; r.F = b.y; r.G = b.z; r.H = b.w; r.I = c.x; r.J = c.y;
  return r;
}

// Hit finalisation (C/raycast.c:461-469, 481-484, 500-523): point and normal of the winner.
RT_DEV void finish_hit(const DevScene &sc, v3 o, v3 d, float t, int idx, v3 &P, v3 &N) {
  P = mk3(add32(o.x, mul32(d.x, t)), add32(o.y, mul32(d.y, t)), add32(o.z, mul32(d.z, t)));
  int2 sl = sc.slot[idx];
  if (sl.x == 0) {
    float4 s = sc.sph[sl.y];
    float inv = (float)__ddiv_rn(1.0, (double)s.w);
    N = norm3(mk3(mul32(sub32(P.x, s.x), inv), mul32(sub32(P.y, s.y), inv), mul32(sub32(P.z, s.z), inv)));
  } else if (sl.x == 1) {
    float4 n = sc.pln[2 * sl.y + 1];
    N = mk3(n.x, n.y, n.z);
  } else {
    quad10 q = load_quad(sc.quad, sl.y);
    double px = P.x, py = P.y, pz = P.z;
    float nx = (float)add64(add64(add64(mul64(mul64(2.0, (double)q.A), px), (double)mul32(q.D, P.y)), (double)mul32(q.E, P.z)), (double)q.G);
    float ny = (float)add64(add64(add64(mul64(mul64(2.0, (double)q.B), py), (double)mul32(q.D, P.x)), (double)mul32(q.F, P.z)), (double)q.H);
    float nz = (float)add64(add64(add64(mul64(mul64(2.0, (double)q.C), pz), (double)mul32(q.E, P.x)), (double)mul32(q.F, P.y)), (double)q.I);
    N = norm3(mk3(nx, ny, nz));
    if (dot3(N, d) > 0.0f) N = mk3(mul32(N.x, -1.0f), mul32(N.y, -1.0f), mul32(N.z, -1.0f));
  }
}

// ---------------------------------------------------------------------------------
// k_primary: one thread per pixel of the tile (grid-stride).
// ---------------------------------------------------------------------------------
template <bool ANY>
RT_DEV bool bvh_trace(const DevScene &sc, const v3 P, const v3 d, const int obj_sph, float &best_t, int &best_idx,
                      unsigned &n_cand, unsigned &n_tests);
template <bool ANY, bool ZRULE>
__device__ __noinline__ bool qbvh_trace(const DevScene &sc, const v3 P, const v3 d, const int obj, float &best_t, int &best_idx,
                                        unsigned &n_cand, unsigned &n_tests);
template <bool ZRULE>
RT_DEV bool quadric_reported(const DevScene &sc, const v3 P, const v3 d, const int kk, float &t, unsigned &n_cand);
RT_DEV void refl_test_sphere(const DevScene &sc, const v3 P, const v3 dj, const int ss, float &best_t, int &best_idx,
                             unsigned &n_cand);
#define RT_PRIM_PX_BRUTE 4   // pixels (= primary rays) per thread: one sphere load feeds four reject tests
template <bool ACCEL>
__global__ void __launch_bounds__(256)
k_primary(DevScene sc, TileGeom tg, float4 *__restrict__ acc, float4 *__restrict__ c0,
          HitQueue q, int nr_max, unsigned long long *__restrict__ counters) {
  constexpr int RT_PRIM_PX = ACCEL ? 1 : RT_PRIM_PX_BRUTE;   // the hierarchy walk gains nothing from sharing sphere loads
  const float pw = __fdiv_rn(sc.cam_w, (float)tg.width);    // C/raycast.c:110
  const float ph = __fdiv_rn(sc.cam_h, (float)tg.height);   // C/raycast.c:109
  const double half_w = __ddiv_rn((double)sc.cam_w, 2.0), half_h = __ddiv_rn((double)sc.cam_h, 2.0);
  unsigned long long n_hits = 0, n_sph = 0, n_pln = 0, n_quad = 0, n_cand = 0, n_ballsum = 0;
  const int stride = gridDim.x * blockDim.x * RT_PRIM_PX;
  const v3 o = mk3(0.0f, 0.0f, 0.0f);
  // block-uniform loop control (the per-thread offset is added inside): the warp stays provably converged, which the
  // full-mask collectives of queue_push and the uniform-register sphere operands both need
  for (int base0 = blockIdx.x * blockDim.x * RT_PRIM_PX; base0 < tg.n_px; base0 += stride) {
    const int base = base0 + threadIdx.x;
    v3 d[RT_PRIM_PX];
    float best_t[RT_PRIM_PX];
    int best_idx[RT_PRIM_PX];
    unsigned live = 0;
    #pragma unroll
    for (int k = 0; k < RT_PRIM_PX; k++) {
      const int i = base + k * blockDim.x;
      best_t[k] = RT_INF; best_idx[k] = -1;
      d[k] = mk3(0.f, 0.f, 0.f);
      if (i < tg.n_px) {
        live |= 1u << k;
        long long g = tg.first_px + i;
        int lr = (int)(g / tg.width), x = (int)(g - (long long)lr * tg.width);
        int y = local_to_image_row(lr, tg.rank, tg.nranks, tg.band_h);
        // C/raycast.c:115-118 (double expression, float store, normalise)
        d[k].x = (float)add64(sub64(0.0, half_w), mul64((double)pw, add64((double)x, 0.5)));
        d[k].y = (float)sub64(add64(0.0, half_h), mul64((double)ph, add64((double)y, 0.5)));
        d[k].z = -1.0f;
        d[k] = norm3(d[k]);
      }
    }
    const int n_live = __popc(live);
    if constexpr (ACCEL) {
      // spheres through the bounding-sphere hierarchy (same literal test for whatever it lets through)
      #pragma unroll 1
      for (int k = 0; k < RT_PRIM_PX; k++) {
        if (!((live >> k) & 1u)) continue;
        unsigned nt = 0, nc = 0;
        const int cell = sc.cam_start != nullptr ? lb_cell(d[k].x, d[k].y, d[k].z) : -1;
        if (cell >= 0) {
          // every primary ray leaves (0,0,0): the camera's own cube map lists what its cell can reach
          const unsigned k1 = sc.cam_start[cell + 1];
          for (unsigned q = sc.cam_start[cell]; q < k1; q++) { nt++; refl_test_sphere(sc, o, d[k], sc.cam_items[q], best_t[k], best_idx[k], nc); }
        } else {
          bvh_trace<false>(sc, o, d[k], -1, best_t[k], best_idx[k], nc, nt);
        }
        n_cand += nc; n_sph += nt;
      }
    } else {
      // spheres: the origin is exactly (0,0,0), so tmp' = -centre and the filter constant k
      // are per-sphere constants prepared on the host.  Operands come from the constant bank (c_sph, c_prim:
      // uniform registers, see rt_scene.cuh); e = d . C is the negative of the chain over tmp' = -C, bit for bit,
      // and only e^2 is used.  Spheres beyond the bank (sc.n_csph) take the same test from global memory.
      auto resolve_pair = [&](const int s, const float m0, const float m1) {
        #pragma unroll 1
        for (int u = 0; u < 2; u++) {
          if ((u ? m1 : m0) < 0.0f) continue;
          const float4 hs = sc.sph_prim[s + u];
          const float4 sp = sc.sph[s + u];
          const int idx = sc.sph_idx[s + u];
          #pragma unroll
          for (int k = 0; k < RT_PRIM_PX; k++) {
            const float e = __fmaf_rn(d[k].z, hs.z, __fmaf_rn(d[k].y, hs.y, d[k].x * hs.x));
            if (__fmaf_rn(e, e, hs.w) >= 0.0f && ((live >> k) & 1u)) {
              n_cand++;
              float t;
              if (sphere_t_literal(o, d[k], sp, t) && better_hit(t, idx, best_t[k], best_idx[k])) { best_t[k] = t; best_idx[k] = idx; }
            }
          }
        }
      };
      const int n_c = min(sc.n_spheres, sc.n_csph);
      #pragma unroll 1
      for (int s = 0; s < n_c; s += 2) {
        const float4 g0 = c_sph[s], g1 = c_sph[s + 1];
        const float k0 = c_prim[s], k1 = c_prim[s + 1];
        float m0 = -1.0f, m1 = -1.0f;
        #pragma unroll
        for (int k = 0; k < RT_PRIM_PX; k++) {
          const float e0 = __fmaf_rn(d[k].z, g0.z, __fmaf_rn(d[k].y, g0.y, d[k].x * g0.x));
          const float e1 = __fmaf_rn(d[k].z, g1.z, __fmaf_rn(d[k].y, g1.y, d[k].x * g1.x));
          m0 = fmaxf(m0, __fmaf_rn(e0, e0, k0));
          m1 = fmaxf(m1, __fmaf_rn(e1, e1, k1));
        }
        if (fmaxf(m0, m1) >= 0.0f) resolve_pair(s, m0, m1);
      }
      for (int s = n_c; s < sc.n_spheres; s += 2) {
        const float4 h0 = sc.sph_prim[s], h1 = sc.sph_prim[s + 1];
        float m0 = -1.0f, m1 = -1.0f;
        #pragma unroll
        for (int k = 0; k < RT_PRIM_PX; k++) {
          const float e0 = __fmaf_rn(d[k].z, h0.z, __fmaf_rn(d[k].y, h0.y, d[k].x * h0.x));
          const float e1 = __fmaf_rn(d[k].z, h1.z, __fmaf_rn(d[k].y, h1.y, d[k].x * h1.x));
          m0 = fmaxf(m0, __fmaf_rn(e0, e0, h0.w));
          m1 = fmaxf(m1, __fmaf_rn(e1, e1, h1.w));
        }
        if (fmaxf(m0, m1) >= 0.0f) resolve_pair(s, m0, m1);
      }
      n_sph += (unsigned long long)sc.n_spheres * n_live;
    }
    for (int p = 0; p < sc.n_planes; p++) {
      float4 pp = sc.pln[2 * p], pn = sc.pln[2 * p + 1];
      v3 n = mk3(pn.x, pn.y, pn.z);
      float num = dot3(sub3(o, mk3(pp.x, pp.y, pp.z)), n);
      int idx = sc.pln_idx[p];
      #pragma unroll
      for (int k = 0; k < RT_PRIM_PX; k++) {
        float t;
        if (((live >> k) & 1u) && plane_t(num, d[k], n, t) && better_hit(t, idx, best_t[k], best_idx[k])) { best_t[k] = t; best_idx[k] = idx; }
      }
    }
    n_pln += (unsigned long long)sc.n_planes * n_live;
    // quadrics: the origin is (0,0,0), so the filter's origin-only part is a per-quadric
    // constant prepared on the host (quad_prim); survivors take the reference's expression tree
    int kq0 = 0;
    if constexpr (ACCEL) {
      // indexed ellipsoids through their hierarchy; the rest below
      if (sc.n_qidx > 0) {
        kq0 = sc.n_qidx;
        #pragma unroll 1
        for (int k = 0; k < RT_PRIM_PX; k++) {
          if (!((live >> k) & 1u)) continue;
          unsigned nt = 0, nc = 0;
          const int cell = sc.camq_start != nullptr ? lb_cell(d[k].x, d[k].y, d[k].z) : -1;
          if (cell >= 0) {
            const unsigned k1 = sc.camq_start[cell + 1];
            for (unsigned q = sc.camq_start[cell]; q < k1; q++) {
              const int kk = sc.camq_items[q];
              float t;
              nt++;
              if (quadric_reported<false>(sc, o, d[k], kk, t, nc) && better_hit(t, sc.quad_idx[kk], best_t[k], best_idx[k])) { best_t[k] = t; best_idx[k] = sc.quad_idx[kk]; }
            }
          } else {
            qbvh_trace<false, false>(sc, o, d[k], -1, best_t[k], best_idx[k], nc, nt);
          }
          n_cand += nc; n_quad += nt;
        }
      }
    }
    unsigned long long n_ball = 0;
    for (int kq = kq0; kq < sc.n_quadrics; kq++) {
      // bounding-ball pre-test (see "Quadrics" at k_shade): the origin is (0,0,0), so T = |c|^2, the radius and kappa are
      // per-quadric constants prepared on the host: {cx, cy, cz, -kappa}; -kappa = +inf for a quadric without a ball
      {
        const float4 hb = sc.quad_ball_prim[kq];
        float mb = -1.0f;
        #pragma unroll
        for (int k = 0; k < RT_PRIM_PX; k++) {
          const float e = __fmaf_rn(d[k].z, hb.z, __fmaf_rn(d[k].y, hb.y, d[k].x * hb.x));
          mb = fmaxf(mb, __fmaf_rn(e, e, hb.w));
        }
        n_ball += (unsigned)n_live;
        if (!(mb >= 0.0f)) continue;
      }
      n_quad += (unsigned)n_live;
      const float4 p0 = sc.quad_prim[3 * kq], p1 = sc.quad_prim[3 * kq + 1], p2 = sc.quad_prim[3 * kq + 2];
      QuadFilter f;
      f.As = p0.x; f.Bs = p0.y; f.Cs = p0.z; f.Ds = p0.w; f.Es = p1.x; f.Fs = p1.y; f.m0 = p1.z; f.eb = p1.w;
      f.gx = p2.x; f.gy = p2.y; f.gz = p2.z;
      float v[RT_PRIM_PX];
      float vmax = -1.0f;
      #pragma unroll
      for (int k = 0; k < RT_PRIM_PX; k++) { v[k] = quad_filter_v(f, d[k]); vmax = fmaxf(vmax, v[k]); }
      if (vmax >= 0.0f) {
        const quad10 qq = load_quad(sc.quad, kq);
        const float c_q = (float)quadric_cq(o, qq);
        const int idx = sc.quad_idx[kq];
        #pragma unroll 1
        for (int k = 0; k < RT_PRIM_PX; k++) {
          float t;   // skip_index == -1: no z rule on primary rays
          if (v[k] >= 0.0f && ((live >> k) & 1u)) {
            n_cand++;
            if (quadric_t(o, d[k], qq, c_q, t) && better_hit(t, idx, best_t[k], best_idx[k])) { best_t[k] = t; best_idx[k] = idx; }
          }
        }
      }
    }
    n_ballsum += n_ball;
    #pragma unroll
    for (int k = 0; k < RT_PRIM_PX; k++) {
      const int i = base + k * blockDim.x;
      const bool hit = i < tg.n_px && best_idx[k] >= 0;
      v3 P = o, N = o;
      int cls = 0;
      if (i < tg.n_px) {
        acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!hit) c0[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (hit) {
        n_hits++;
        finish_hit(sc, o, d[k], best_t[k], best_idx[k], P, N);
        cls = hit_class(sc, P, N, best_idx[k], sc.max_depth > 1, nr_max);
      }
      queue_push(q, hit, make_float4(P.x, P.y, P.z, __int_as_float(i)), make_float4(N.x, N.y, N.z, __int_as_float(best_idx[k])),
                 make_float4(d[k].x, d[k].y, d[k].z, 1.0f), cls);
    }
  }
  // counters: one atomic per warp per counter
  #pragma unroll
  for (int off = 16; off; off >>= 1) {
    n_hits += __shfl_xor_sync(0xffffffffu, n_hits, off);
    n_sph += __shfl_xor_sync(0xffffffffu, n_sph, off);
    n_pln += __shfl_xor_sync(0xffffffffu, n_pln, off);
    n_quad += __shfl_xor_sync(0xffffffffu, n_quad, off);
    n_cand += __shfl_xor_sync(0xffffffffu, n_cand, off);
    n_ballsum += __shfl_xor_sync(0xffffffffu, n_ballsum, off);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&counters[CT_BALL_D], n_ballsum);
    atomicAdd(&counters[CT_PRIMARY_HITS], n_hits);
    atomicAdd(&counters[CT_SPH_D], n_sph);
    atomicAdd(&counters[CT_PLN_D], n_pln);
    atomicAdd(&counters[CT_QUAD_D], n_quad);
    atomicAdd(&counters[CT_QUAD_O], n_quad);
    atomicAdd(&counters[CT_PLN_O], n_pln);
    atomicAdd(&counters[CT_CANDIDATES], n_cand);
  }
}

// ---------------------------------------------------------------------------------
// k_shade<LB>: one thread per hit point of wavefront level `level` (grid-stride).
// Ray slot 0 is the reflection ray, slots 1..LB the shadow rays of one pass of LB
// lights.  All slots share origin P and skip index, so per primitive the origin-only
// part (sphere: tmp, c; plane: num; quadric: c_q) is computed once.
// ---------------------------------------------------------------------------------
// Resolve stage of the sphere scan, shared by all scan variants (one copy: instruction-cache room is what
// limits how many variants can exist): every survivor of the chunk against every ray still alive, by the
// filter recheck, the sign analysis, or the reference's literal root.
// Survivors arrive as bit masks in registers (RT_CHUNK = 128 spheres = 4 words).  Round 1 queued their indices in
// shared memory with a branch-free store of the loop counter; that store needs the counter in a per-thread
// register, and ptxas then derives the sphere address from the same register and loads the operands per thread
// -- no uniform registers, three-register FFMAs (tools: cuobjdump -sass | grep "FFMA.*UR").
__device__ __noinline__ void resolve_spheres(const float (*dsm)[128], const int tid, const v3 P,
                                             const DevScene &sc, const int cs, const int nw,
                                             const unsigned w0, const unsigned w1, const unsigned w2, const unsigned w3,
                                             const int obj, const bool r_on,
                                             unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand) {
  // survivors of the chunk: one bit per sphere, 32 spheres per word, the first sphere of a word in bit 31 (the scans shift
  // their bits in from the right); w0 is the LAST word scanned (word nw-1), w1 the one before, ...
  for (int k = nw - 1; k >= 0; k--)
  for (unsigned sm = k == 0 ? w0 : k == 1 ? w1 : k == 2 ? w2 : w3; sm; sm &= ~(0x80000000u >> __clz(sm))) {
    const int ss = cs + ((nw - 1 - k) << 5) + __clz(sm);
    if (ss >= sc.n_spheres) continue;                          // a pad entry can only get here with a NaN margin (|P|^2 overflow)
    const int idx = sc.sph_idx[ss];
    if (idx == obj) continue;                                  // C/raycast.c:450
    const float4 hh = sc.sph_hot[ss];
    const float xs = P.x - hh.x, ys = P.y - hh.y, zs = P.z - hh.z;
    const float Ts = __fmaf_rn(zs, zs, __fmaf_rn(ys, ys, xs * xs));
    const float cp = Ts - hh.w, marg = RT_FILTER_EPS * (Ts + hh.w);
    const float lo = cp - marg, hi = cp + marg;
    for (unsigned pm = alive; pm; pm &= pm - 1) {
      const int j = __ffs(pm) - 1;
      const v3 dj = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
      const float e = __fmaf_rn(dj.z, zs, __fmaf_rn(dj.y, ys, dj.x * xs));
      const float e2 = e * e;
      if (!(e2 >= lo)) continue;                               // the filter rejects this ray
      // Sign analysis: with the discriminant certainly >= 0 (e2 >= hi, or the
      // origin certainly inside) and c certainly away from 0 relative to e2, the reference's
      // root selection (C/raycast.c:593-597) yields t > 0 exactly when the origin is inside,
      // or outside with the sphere ahead (dt < 0).  A shadow ray needs no root then;
      // whatever is uncertain takes the literal path.
      const float guard = 1.52587890625e-05f * e2;             // 2^-16 e^2
      const bool inside = cp <= -(marg + guard);
      const bool outside = (cp >= marg + guard) && (e2 >= hi);
      const bool is_refl = r_on && j == 0;
      const bool need_lit = is_refl ? !(outside && e > 0.0f) : !(inside || outside);
      bool occl = !is_refl && (inside || (outside && e < 0.0f));
      if (need_lit) {
        n_cand++;
        float t;
        if (sphere_t_literal(P, dj, sc.sph[ss], t)) {
          if (is_refl) { if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; } }
          else occl = (t > 0.0f) && (t < RT_INF);
        }
      }
      if (occl) alive &= ~(1u << j);
    }
  }
}

// Sphere scan of one chunk with NRV ray slots held in registers (loaded from the shared-memory
// mirror), operands from GLOBAL memory: the round-1 loop, kept for the spheres that do not fit the
// constant bank (scan_chunk_c below is the production path).  Conservative reject filter (DESIGN.md
// "sphere filter"): with tmp' = P - C, e_j = fma-chain dot(d_j, tmp'), T' = |tmp'|^2,
// kappa = (T' - r^2) - 2^-18 (T' + r^2):
//   max_j e_j^2 < kappa  proves the reference's discriminant (C/raycast.c:587) negative for
//   every ray leaving P, so none of them hits this sphere.
// Survivors are recorded branch-free as bits (32 spheres per word, see resolve_spheres) and resolved after
// the chunk, so the scan never diverges.  The arrays are padded to whole words with entries that never pass
// (r^2 = -inf => kappa = +inf), plus one prefetch group.
// scan_balls is the loop itself over any table of {centre, r^2} with margin eps: the spheres (eps = 2^-18), and the
// bounding balls of the quadrics (DevScene::quad_ball), whose survivors go to resolve_quadrics.
// QB = false: table entries {centre, r^2}, kappa = (T - r^2) - eps (T + r^2) (the sphere filter).
// QB = true: bounding balls of quadrics, entries {centre, a}: the ball's radius grows with the distance of the origin,
// R = a + eps sqrt(T) (rt_abi.cu, "Bounding balls"), kappa = T (1 - 2^-16) - R |R|; a = +inf (no certified ball) gives
// kappa = -inf (always passes), a = -inf (pads) +inf (never passes).  sqrt(T) = T rsqrt(T) rounded up by 2^-20.
template <int NRV, bool QB>
RT_DEV uint4 scan_balls(const float (*dsm)[128], const int tid, const v3 P, const float4 *__restrict__ table, const float eps,
                        const int cs, const int n) {
  const int nw = (n + 31) >> 5;
  const float4 *hp = table + cs;
  v3 d[NRV];
  #pragma unroll
  for (int j = 0; j < NRV; j++) d[j] = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
  unsigned w0 = 0, w1 = 0, w2 = 0, w3 = 0;
  float4 a0 = hp[0], a1 = hp[1];
  #pragma unroll 1
  for (int w = 0; w < nw; w++) {
    unsigned wc = 0;
    const int cnt = (min(32, n - (w << 5)) + 1) & ~1;       // entries of this word, whole trips (a ragged last word: small scenes)
    #pragma unroll 1
    for (int s = 0; s < cnt; s += 2) {
      hp += 2;
      const float4 g0 = a0, g1 = a1;
      a0 = hp[0]; a1 = hp[1];                                // next trip's operands (spare entries at the end)
      const float x0 = P.x - g0.x, y0 = P.y - g0.y, z0 = P.z - g0.z;
      const float x1 = P.x - g1.x, y1 = P.y - g1.y, z1 = P.z - g1.z;
      const float T0 = __fmaf_rn(z0, z0, __fmaf_rn(y0, y0, x0 * x0));
      const float T1 = __fmaf_rn(z1, z1, __fmaf_rn(y1, y1, x1 * x1));
      float ka0, ka1;
      if constexpr (QB) {
        const float R0 = __fmaf_rn(eps, T0 * rsqrtf(T0) * 1.000001f, g0.w), R1 = __fmaf_rn(eps, T1 * rsqrtf(T1) * 1.000001f, g1.w);
        ka0 = __fmaf_rn(T0, 0.9999847412109375f, -(R0 * fabsf(R0)));
        ka1 = __fmaf_rn(T1, 0.9999847412109375f, -(R1 * fabsf(R1)));
      } else {
        ka0 = __fmaf_rn(-eps, T0 + g0.w, T0 - g0.w);
        ka1 = __fmaf_rn(-eps, T1 + g1.w, T1 - g1.w);
      }
      float mx0 = 0.0f, mx1 = 0.0f;
      #pragma unroll
      for (int j = 0; j < NRV; j++) {
        mx0 = fmaxf(mx0, fabsf(__fmaf_rn(d[j].z, z0, __fmaf_rn(d[j].y, y0, d[j].x * x0))));
        mx1 = fmaxf(mx1, fabsf(__fmaf_rn(d[j].z, z1, __fmaf_rn(d[j].y, y1, d[j].x * x1))));
      }
      wc = __funnelshift_l(__float_as_uint(__fmaf_rn(mx0, mx0, -ka0)), wc, 1);   // sign bit set = rejected (see scan_chunk_c)
      wc = __funnelshift_l(__float_as_uint(__fmaf_rn(mx1, mx1, -ka1)), wc, 1);
    }
    w3 = w2; w2 = w1; w1 = w0; w0 = (~wc) << (32 - cnt);     // first entry of the word in bit 31, entries never scanned = 0
  }
  return make_uint4(w0, w1, w2, w3);
}
template <int NRV>
RT_DEV void scan_chunk_g(const float (*dsm)[128], const int tid, const v3 P,
                         const DevScene &sc, const int cs, const int n, const int obj, const bool r_on,
                         unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand) {
  const int nw = (n + 31) >> 5;
  const uint4 w = scan_balls<NRV, false>(dsm, tid, P, sc.sph_hot, RT_FILTER_EPS, cs, n);
  resolve_spheres(dsm, tid, P, sc, cs, nw, w.x, w.y, w.z, w.w, obj, r_on, alive, best_t, best_idx, n_cand);
}

// The same scan with the sphere operands in the constant bank (rt_scene.cuh, c_sph): the production path.
// Per ray the chunk-invariant q_j = d_j . P is formed once; per (ray, sphere)
//     e'_j = fma(-dz, cz, fma(-dy, cy, fma(-dx, cx, q_j)))  ~  d_j . (P - C)
// is three FFMAs whose middle operand is a uniform register, so each reads two register-file operands and issues at
// full rate (a three-register FFMA does not: 0.61 per clock on B200).  tmp = P - C is only needed for T' now.
// Conservativeness (u = 2^-24; dt, a, c as in the contract above; E* = d.(P - C) in exact arithmetic):
//   |dt - E*| <= 4.2u |tmp|   (rounding of tmp, three roundings of the reference's dot product),
//   |e' - E*| <= 6.1u (|P| + |C|)   (six roundings, every intermediate bounded by |d| (|P| + |C|)),
//   dt^2 <= e'^2 + 2 |e'| delta + delta^2,  delta = 6.1u (|P| + |C|) + 4.2u |tmp|,  |e'| <= 1.001 |tmp| + delta,
//   2.01 |tmp| delta <= 14.7u T + 6.15u (|P| + |C|)^2 <= 14.7u T + 12.3u (|P|^2 + |C|^2),
// and the reference rejects whenever dt^2 < (T' - r^2) - 20u (T' + r^2) (the contract's budget without its own 15u for e).
// Hence  e'^2 < kappa' = (T' - r^2) - 2^-18 (T' + r^2) - 2^-19 (|P|^2 + |C|^2)  proves the reference's miss:
// 64u covers 14.8u + 20u and the roundings of kappa' and m^2, 32u covers 12.4u and the roundings of |P|^2, |C|^2
// (both rounded up: rho on the host, RT_P2_EPS here).  kappa' = fma(T', 1 - 2^-18, -(rho + p2s)).  The extra term makes
// the filter let through rays within a relative 2^-19 (|P|^2 + |C|^2) / T of tangency -- a few per mille more survivors
// for content tens of units from the origin -- and survivors are decided by resolve_spheres as before (tight filter
// on tmp, sign analysis, literal root), so the image cannot change.
// Spheres per loop trip: the widest variants take one (the trip is ~75 instructions), the narrow ones four.
template <int NRV>
RT_DEV void scan_chunk_c(const float (*dsm)[128], const int tid, const v3 P, const float p2s,
                         const DevScene &sc, const int cs, const int n, const int obj, const bool r_on,
                         unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand) {
  float nx[NRV], ny[NRV], nz[NRV], q[NRV];
  #pragma unroll
  for (int j = 0; j < NRV; j++) {
    const float dx = dsm[3 * j][tid], dy = dsm[3 * j + 1][tid], dz = dsm[3 * j + 2][tid];
    nx[j] = -dx; ny[j] = -dy; nz[j] = -dz;
    q[j] = __fmaf_rn(dz, P.z, __fmaf_rn(dy, P.y, dx * P.x));
  }
  constexpr int U = NRV >= RT_TRIP1_MIN ? 1 : NRV >= RT_TRIP2_MIN ? 2 : 4;
  unsigned w0 = 0, w1 = 0, w2 = 0, w3 = 0;
  const int nw = (n + 31) >> 5;                              // whole words: the bank is padded with entries that never pass
  float4 gn[U];                                              // operands of the next trip, in flight during this one
  #pragma unroll
  for (int u = 0; u < U; u++) gn[u] = c_sph[cs + u];
#if RT_PREFETCH2
  float4 gn2[U];                                             // ... and of the one after (the constant load's latency is ~1 trip)
  #pragma unroll
  for (int u = 0; u < U; u++) gn2[u] = c_sph[cs + U + u];
#endif
  #pragma unroll 1
  for (int w = 0; w < nw; w++) {
    unsigned wc = 0;
    const int cnt = (min(32, n - (w << 5)) + U - 1) & ~(U - 1);   // entries of this word, whole trips (a ragged last word: small scenes)
    #pragma unroll 1
    for (int s = 0; s < cnt; s += U) {
      float4 gc[U];
      #pragma unroll
#if RT_PREFETCH2
      for (int u = 0; u < U; u++) { gc[u] = gn[u]; gn[u] = gn2[u]; gn2[u] = c_sph[cs + (w << 5) + s + 2 * U + u]; }
#else
      for (int u = 0; u < U; u++) { gc[u] = gn[u]; gn[u] = c_sph[cs + (w << 5) + s + U + u]; }   // warp-uniform index: uniform registers
#endif
      #pragma unroll
      for (int u = 0; u < U; u++) {
        const float4 g = gc[u];
        const float x = P.x - g.x, y = P.y - g.y, z = P.z - g.z;
        const float T = __fmaf_rn(z, z, __fmaf_rn(y, y, x * x));
        const float ka = __fmaf_rn(T, 1.0f - RT_FILTER_EPS, -(g.w + p2s));
        float m = 0.0f;
        #pragma unroll
        for (int j = 0; j < NRV; j++)
          m = fmaxf(m, fabsf(__fmaf_rn(nz[j], g.z, __fmaf_rn(ny[j], g.y, __fmaf_rn(nx[j], g.x, q[j])))));
        // survivor bit: the SIGN of m^2 - kappa' (one FFMA, exact sign) shifted into the word (one SHF) -- first sphere of
        // the word ends up in bit 31, a set bit means REJECTED; a NaN kappa' gives a clear bit = survivor, resolved exactly
        wc = __funnelshift_l(__float_as_uint(__fmaf_rn(m, m, -ka)), wc, 1);
      }
    }
    w3 = w2; w2 = w1; w1 = w0; w0 = (~wc) << (32 - cnt);       // first sphere of the word in bit 31, spheres never scanned = 0
  }
  resolve_spheres(dsm, tid, P, sc, cs, nw, w0, w1, w2, w3, obj, r_on, alive, best_t, best_idx, n_cand);
}

// Quadrics.  Every ray still meets every quadric, in two stages:
//  (1) a conservative test against the quadric's BOUNDING BALL (DevScene::quad_ball, scan_balls): for an ellipsoid the
//      host can certify (rt_qindex.cuh, qindex_classify: positive definite, well conditioned, real surface) a line that
//      passes the centre farther than r_item + rho |P - c| has a negative discriminant IN THE REFERENCE'S FLOAT
//      ARITHMETIC (claim (1) there), so the quadric cannot be reported for any ray of the hit point; every other
//      quadric (cylinders, cones, paraboloids, flat or degenerate ones) has an infinite ball and always passes.
//      Cost: the sphere filter's 8 + 3 per ray instead of the quadric filter's 45 + 14 per ray.
//  (2) survivors, per thread: the FP32 quadric reject filter (rt_math.cuh) per ray still alive, then the reference's
//      expression tree (quadric_cq / quadric_t, mixed float/double) and the z rule of C/raycast.c:492-494.
__device__ __noinline__ void resolve_quadrics(const float (*dsm)[128], const int tid, const v3 P,
                                              const DevScene &sc, const int cs, const int nw,
                                              const unsigned w0, const unsigned w1, const unsigned w2, const unsigned w3,
                                              const int obj, const bool r_on,
                                              unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand, unsigned &n_setup, unsigned &n_filt) {
  // no ray left to test (every shadow ray already occluded, or a stand-in lane without a hit point): the quadrics that
  // passed their ball -- all of them when they have no certified ball -- need no filter set-up for this thread
  if (!alive) return;
  for (int k = nw - 1; k >= 0; k--)
  for (unsigned sm = k == 0 ? w0 : k == 1 ? w1 : k == 2 ? w2 : w3; sm; sm &= ~(0x80000000u >> __clz(sm))) {
    const int kk = cs + ((nw - 1 - k) << 5) + __clz(sm);
    if (kk >= sc.n_quadrics) continue;
    const int idx = sc.quad_idx[kk];
    if (idx == obj) continue;                                  // C/raycast.c:450
    const QuadFilter f = quad_filter_setup(sc.quad_hot[3 * kk], sc.quad_hot[3 * kk + 1], sc.quad_hot[3 * kk + 2], P.x, P.y, P.z);
    n_setup++;
    unsigned pm = 0;
    for (unsigned m = alive; m; m &= m - 1) {
      const int j = __ffs(m) - 1;
      const v3 dj = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
      pm |= (unsigned)(quad_filter_v(f, dj) >= 0.0f) << j;
      n_filt++;
    }
    if (!pm) continue;
    const quad10 qq = load_quad(sc.quad, kk);
    const float c_q = (float)quadric_cq(P, qq);
    for (; pm; pm &= pm - 1) {
      const int j = __ffs(pm) - 1;
      const v3 dj = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
      float t;
      n_cand++;
      if (quadric_t(P, dj, qq, c_q, t)) {
        if (add32(P.z, mul32(t, dj.z)) < P.z) continue;       // C/raycast.c:492-494 (skip_index != -1 here)
        if (r_on && j == 0) { if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; } }
        else if (t > 0.0f && t < RT_INF) alive &= ~(1u << j);
      }
    }
  }
}

template <int NRV>
RT_DEV void scan_quadrics(const float (*dsm)[128], const int tid, const v3 P,
                          const DevScene &sc, const int cs, const int n, const int obj, const bool r_on,
                          unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand, unsigned &n_setup, unsigned &n_filt) {
  const int nw = (n + 31) >> 5;
  const uint4 w = scan_balls<NRV, true>(dsm, tid, P, sc.quad_ball, sc.quad_ball_eps, cs, n);
  resolve_quadrics(dsm, tid, P, sc, cs, nw, w.x, w.y, w.z, w.w, obj, r_on, alive, best_t, best_idx, n_cand, n_setup, n_filt);
}

// One shadow ray (origin P, direction dj) against sphere slot ss: does the reference report a hit with
// 0 < t < inf?  Same decision chain as the resolve stage of scan_chunk: conservative filter, sign
// analysis where the margins allow it, the reference's literal root otherwise.
RT_DEV bool shadow_hits_sphere(const DevScene &sc, const v3 P, const v3 dj, const int ss, const float4 hh, unsigned &n_cand) {
  const float xs = P.x - hh.x, ys = P.y - hh.y, zs = P.z - hh.z;
  const float Ts = __fmaf_rn(zs, zs, __fmaf_rn(ys, ys, xs * xs));
  const float cp = Ts - hh.w, marg = RT_FILTER_EPS * (Ts + hh.w);
  const float lo = cp - marg, hi = cp + marg;
  const float e = __fmaf_rn(dj.z, zs, __fmaf_rn(dj.y, ys, dj.x * xs));
  const float e2 = e * e;
  if (!(e2 >= lo)) return false;
  const float guard = 1.52587890625e-05f * e2;
  const bool inside = cp <= -(marg + guard);
  const bool outside = (cp >= marg + guard) && (e2 >= hi);
  if (inside || outside) return inside || e < 0.0f;
  n_cand++;
  float t;
  return sphere_t_literal(P, dj, sc.sph[ss], t) && (t > 0.0f) && (t < RT_INF);
}

// One reflection ray against sphere slot ss (nearest hit): the decision chain of scan_chunk's resolve stage.
RT_DEV void refl_test_sphere(const DevScene &sc, const v3 P, const v3 dj, const int ss, float &best_t, int &best_idx,
                             unsigned &n_cand) {
  const float4 hh = sc.sph_hot[ss];
  const float xs = P.x - hh.x, ys = P.y - hh.y, zs = P.z - hh.z;
  const float Ts = __fmaf_rn(zs, zs, __fmaf_rn(ys, ys, xs * xs));
  const float cp = Ts - hh.w, marg = RT_FILTER_EPS * (Ts + hh.w);
  const float lo = cp - marg, hi = cp + marg;
  const float e = __fmaf_rn(dj.z, zs, __fmaf_rn(dj.y, ys, dj.x * xs));
  const float e2 = e * e;
  if (!(e2 >= lo)) return;
  const float guard = 1.52587890625e-05f * e2;
  const bool outside = (cp >= marg + guard) && (e2 >= hi);
  if (outside && e > 0.0f) return;                             // both roots behind the origin
  n_cand++;
  float t;
  const int idx = sc.sph_idx[ss];
  if (sphere_t_literal(P, dj, sc.sph[ss], t) && better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; }
}

// Traversal of the bounding-sphere hierarchy for one ray.  A node is entered unless the ray certainly
// misses every member in the REFERENCE's arithmetic: the reference can accept sphere (C, r) only if the
// line passes within sqrt(r^2 + 2^-18 (T + r^2)) <= r (1 + 2^-19) + 2^-9 sqrt(T) of C (reject-filter contract:
// its error budget, 35u (T + r^2), plus the roundings of this restatement, 59u in all; T = |P - C|^2 <=
// (|P - Cn| + Rn)^2 for a member of node (Cn, Rn)), so the node sphere is inflated by that much plus the
// float slack of this test itself; nodes wholly behind the origin (with the larger 2^-6 margin the
// "both roots negative" argument needs) and, for nearest hit, wholly beyond the best t so far are skipped.
// obj_sph: sphere slot of the object the ray leaves (-1: none / not a sphere), skipped like C/raycast.c:450.
// ANY = true: shadow query, returns at the first hit.  Returns whether anything was hit (ANY) / unused.
template <bool ANY>
RT_DEV bool bvh_trace(const DevScene &sc, const v3 P, const v3 d, const int obj_sph, float &best_t, int &best_idx,
                      unsigned &n_cand, unsigned &n_tests) {
  // node test: enter?  key: no member can report a t below key / 1.001 (the "beyond" margin taken off already)
  auto enter = [&](const float4 nd, float &key) -> bool {
    const float wx = P.x - nd.x, wy = P.y - nd.y, wz = P.z - nd.z;
    const float Tn = __fmaf_rn(wz, wz, __fmaf_rn(wy, wy, wx * wx));
    const float en = __fmaf_rn(d.z, wz, __fmaf_rn(d.y, wy, d.x * wx));
    const float reach = sqrtf(Tn) + nd.w;                      // >= sqrt(T) of every member
    const float rinf = __fmaf_rn(nd.w, 1.0001f, 0.0029296875f * reach);  // Rn (1 + 1e-4) + (2^-9 + 2^-10) reach
    // line misses the inflated node?  (Tn - en^2 is the squared distance, float slack 1e-5 Tn)
    if (!(Tn - en * en <= __fmaf_rn(rinf, rinf, 1e-5f * Tn))) return false;
    const float rbehind = __fmaf_rn(nd.w, 1.001f, 0.015625f * reach);     // Rn (1 + 1e-3) + 2^-6 reach
    if (en > rbehind) return false;                            // whole node behind the origin
    key = (-en - rinf) - 4e-3f * reach;
    return ANY || !(key > best_t * 1.001f);                    // whole node beyond the best hit
  };
  int stack[32];
  float skey[32];
  int sp = 0, k = 0;
  float key0;
  if (!enter(sc.bvh_node[0], key0)) return false;
  auto pop = [&]() -> bool {
    while (sp > 0) {
      --sp;
      if (ANY || !(skey[sp] > best_t * 1.001f)) { k = stack[sp]; return true; }   // best_t may have come down since the push
    }
    return false;
  };
  // "while-while": walk inner nodes until a leaf turns up, then test the leaf -- the lanes of a warp spend their
  // time in the same one of the two loops
  for (;;) {
    int2 kid = sc.bvh_kids[k];
    while (kid.x >= 0) {
      // both children's spheres come with the parent: test both, go into the nearer one (children were split
      // along axis kid.y >> 28, left = lower coordinates), keep the other for later
      const int axis = kid.y >> 28, right = kid.y & 0x0fffffff;
      const float4 nl = sc.bvh_pair[2 * k], nr = sc.bvh_pair[2 * k + 1];
      float kl = 0.f, kr = 0.f;
      const bool hl = enter(nl, kl), hr = enter(nr, kr);
      const float da = axis == 0 ? d.x : axis == 1 ? d.y : d.z;
      const bool left_first = da >= 0.0f;
      if (hl && hr) {
        stack[sp] = left_first ? right : kid.x;                // depth <= 28 (host), one net entry per level
        skey[sp++] = left_first ? kr : kl;
        k = left_first ? kid.x : right;
      } else if (hl || hr) {
        k = hl ? kid.x : right;
      } else if (!pop()) {
        return false;
      }
      kid = sc.bvh_kids[k];
    }
    const int first = -kid.x - 1;
    for (int i = 0; i < kid.y; i++) {
      const int ss = first + i;                                // leaves are runs of consecutive sphere slots
      n_tests++;
      if (ss == obj_sph) continue;                             // C/raycast.c:450
      if (ANY) { if (shadow_hits_sphere(sc, P, d, ss, sc.sph_hot[ss], n_cand)) return true; }
      else refl_test_sphere(sc, P, d, ss, best_t, best_idx, n_cand);
    }
    if (!pop()) return false;
  }
}

// One ray against quadric slot kk, the chain of resolve_quadrics: reject filter, then the reference's expression
// tree and (ZRULE) the hit-point rule of C/raycast.c:492-494.  Returns whether the reference reports a t; the
// caller applies its own acceptance (0 < t < inf for a shadow ray, better_hit for a nearest-hit search).
template <bool ZRULE>
RT_DEV bool quadric_reported(const DevScene &sc, const v3 P, const v3 d, const int kk, float &t, unsigned &n_cand) {
  const QuadFilter f = quad_filter_setup(sc.quad_hot[3 * kk], sc.quad_hot[3 * kk + 1], sc.quad_hot[3 * kk + 2], P.x, P.y, P.z);
  if (!(quad_filter_v(f, d) >= 0.0f)) return false;
  const quad10 qq = load_quad(sc.quad, kk);
  const float c_q = (float)quadric_cq(P, qq);
  n_cand++;
  if (!quadric_t(P, d, qq, c_q, t)) return false;
  if (ZRULE && add32(P.z, mul32(t, d.z)) < P.z) return false;
  return true;
}

template <bool ANY, bool ZRULE>
__device__ __noinline__ bool qbvh_trace(const DevScene &sc, const v3 P, const v3 d, const int obj, float &best_t, int &best_idx,
                                        unsigned &n_cand, unsigned &n_tests) {
  auto enter = [&](const float4 nd, const float4 mg, float &key) -> bool {
    const QIndexConsts qk = {mg.x, sc.q_behind_rel, mg.y, mg.z, mg.w};
    return qnode_prune(nd, P.x, P.y, P.z, d.x, d.y, d.z, !ANY, best_t, qk, &key) == 0;
  };
  int stack[32];
  float skey[32];
  int sp = 0, k = 0;
  float key0;
  if (!enter(sc.qbvh_node[0], sc.qbvh_marg[0], key0)) return false;
  auto pop = [&]() -> bool {
    while (sp > 0) {
      --sp;
      if (ANY || !(skey[sp] > best_t * 1.001f)) { k = stack[sp]; return true; }
    }
    return false;
  };
  for (;;) {
    int2 kid = sc.qbvh_kids[k];
    while (kid.x >= 0) {
      const int axis = kid.y >> 28, right = kid.y & 0x0fffffff;
      const float4 nl = sc.qbvh_pair[2 * k], nr = sc.qbvh_pair[2 * k + 1];
      const float4 ml = sc.qbvh_mpair[2 * k], mr = sc.qbvh_mpair[2 * k + 1];
      float kl = 0.f, kr = 0.f;
      const bool hl = enter(nl, ml, kl), hr = enter(nr, mr, kr);
      const float da = axis == 0 ? d.x : axis == 1 ? d.y : d.z;
      const bool left_first = da >= 0.0f;
      if (hl && hr) {
        stack[sp] = left_first ? right : kid.x;                // depth <= 28 (host), one net entry per level
        skey[sp++] = left_first ? kr : kl;
        k = left_first ? kid.x : right;
      } else if (hl || hr) {
        k = hl ? kid.x : right;
      } else if (!pop()) {
        return false;
      }
      kid = sc.qbvh_kids[k];
    }
    const int first = -kid.x - 1;
    for (int i = 0; i < kid.y; i++) {
      const int kk = first + i;                                // leaves are runs of consecutive quadric slots
      n_tests++;
      const int idx = sc.quad_idx[kk];
      if (idx == obj) continue;                                // C/raycast.c:450
      float t;
      if (!quadric_reported<ZRULE>(sc, P, d, kk, t, n_cand)) continue;
      if (ANY) { if (t > 0.0f && t < RT_INF) return true; }
      else if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; }
    }
    if (!pop()) return false;
  }
}

// Packs the ray slots still alive to the front of the slot array (directions in dsm, light numbers in lmap) and
// fills the rest of the nv slots of the next scan variant with NaN (they lose every max and pass no test).
// Out of line on purpose: inlined at its call sites inside the chunk loop, this code (per-thread loops around
// shared-memory moves) made ptxas give up on the uniform datapath for the whole kernel -- the scan's sphere
// operands then came through per-thread registers (cuobjdump -sass | grep "FFMA.*UR" showed none).
__device__ __noinline__ void repack_slots(float (*dsm)[128], const int tid, const int sh, const int nv,
                                          unsigned &alive, unsigned long long &lmap) {
  int k = 0;
  unsigned long long nl = 0;
  for (unsigned m = alive; m; m &= m - 1) {
    const int j = __ffs(m) - 1;
    if (k != j) { dsm[3 * k][tid] = dsm[3 * j][tid]; dsm[3 * k + 1][tid] = dsm[3 * j + 1][tid]; dsm[3 * k + 2][tid] = dsm[3 * j + 2][tid]; }
    if (j >= sh) nl |= ((lmap >> (4 * (j - sh))) & 15ull) << (4 * (k - sh));
    k++;
  }
  lmap = nl;
  alive = (1u << k) - 1u;
  for (int j = k; j < nv; j++) { dsm[3 * j][tid] = RT_NAN; dsm[3 * j + 1][tid] = RT_NAN; dsm[3 * j + 2][tid] = RT_NAN; }
}

// Scan variants: the bundle scans are compiled for NR, NR-2, NR-4, ... >= 3 ray slots; a warp runs
// the smallest one that holds the largest number of rays any of its hit points still has alive.
template <int NRV, int NRMIN>
RT_DEV void scan_spheres_v(const int nrv, const float (*dsm)[128], const int tid, const v3 P, const float p2s,
                           const DevScene &sc, const int cs, const int n, const int obj, const bool r_on,
                           unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand) {
  if constexpr (NRV - RT_VSTEP >= NRMIN) {
    if (nrv < NRV) { scan_spheres_v<NRV - RT_VSTEP, NRMIN>(nrv, dsm, tid, P, p2s, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand); return; }
  }
  scan_chunk_c<NRV>(dsm, tid, P, p2s, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand);
}
// Chunks beyond the constant bank (scenes with more than RT_CSPH_CAP spheres): the global-operand loop, one copy per
// variant but out of line -- it only has to be correct.
// These variants come in steps of 4 slots while the caller packs and NaN-fills the slots in steps of RT_VSTEP: the
// slots between the caller's nrv and the compiled width are filled here, or they would carry stale directions (of
// rays already occluded, or of the previous hit point) through the filter and inflate the survivor lists.
template <int NRV, int NRMIN>
__device__ __noinline__ void scan_spheres_g(const int nrv, float (*dsm)[128], const int tid, const v3 P,
                                            const DevScene &sc, const int cs, const int n, const int obj, const bool r_on,
                                            unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand) {
  if constexpr (NRV - 4 >= NRMIN) {
    if (nrv <= NRV - 4) { scan_spheres_g<NRV - 4, NRMIN>(nrv, dsm, tid, P, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand); return; }
  }
  for (int j = nrv; j < NRV; j++) { dsm[3 * j][tid] = RT_NAN; dsm[3 * j + 1][tid] = RT_NAN; dsm[3 * j + 2][tid] = RT_NAN; }
  scan_chunk_g<NRV>(dsm, tid, P, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand);
}
template <int NRV, int NRMIN>
RT_DEV void scan_quadrics_v(const int nrv, const float (*dsm)[128], const int tid, const v3 P,
                            const DevScene &sc, const int cs, const int n, const int obj, const bool r_on,
                            unsigned &alive, float &best_t, int &best_idx, unsigned &n_cand, unsigned &n_setup, unsigned &n_filt) {
  if constexpr (NRV - RT_VSTEP >= NRMIN) {
    if (nrv < NRV) { scan_quadrics_v<NRV - RT_VSTEP, NRMIN>(nrv, dsm, tid, P, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand, n_setup, n_filt); return; }
  }
  scan_quadrics<NRV>(dsm, tid, P, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand, n_setup, n_filt);
}

// QUADS = false: compiled without the quadric sections for scenes that have none (half the code:
// the instruction cache is the scarce resource of this kernel).
// ACCEL = true: the sphere part of every shadow query goes through the light buffer (rt_scene.cuh)
// instead of the bundle scan; the reflection ray still scans every sphere.
template <int LB, bool QUADS, bool ACCEL>
__global__ void __launch_bounds__(128, (ACCEL && !QUADS) ? 2 * RT_SHADE_MINB : (LB <= 8 ? RT_SHADE_MINB_NARROW : RT_SHADE_MINB))   // no bundle scan: half the registers, twice the warps
k_shade(DevScene sc, int level, HitQueue qin, HitQueue qout, float4 *__restrict__ acc,
        float4 *__restrict__ c0, unsigned long long *__restrict__ counters, float4 *__restrict__ lvl) {
  constexpr int NR = LB + 1;
  constexpr int VMIN = NR >= 3 ? NR - RT_VSTEP * ((NR - 3) / RT_VSTEP) : NR;   // smallest compiled variant
  auto pick_variant = [](int live) { const int v = NR - RT_VSTEP * ((NR - live) / RT_VSTEP); return v < VMIN ? VMIN : v; };
  __shared__ float dsm[3 * NR][128];
  const int tid = threadIdx.x;
  // per-thread event counts (32-bit: a thread sees < 2^20 items per launch); the per-type test
  // counters are products of these with the scene's primitive counts, formed at the end
  unsigned n_refl = 0, n_lit = 0, n_hp = 0, n_pass_sum = 0, n_ls = 0, n_cand = 0;
  unsigned d_sph = 0, d_pln = 0, d_quad = 0, d_slots = 0, o_sph = 0, o_quad = 0, d_ball = 0, o_ball = 0;   // direction tests executed: live ray slots x primitives scanned with them
  // Dynamic scheduling: every warp claims 32 consecutive entries of one class list at a time
  // (heaviest class first), so no SM idles while another still holds a backlog, and the hit
  // points a warp shades together need about the same number of ray slots.
  int cstart[RT_NCLS + 1];   // virtual index space: class lists padded to multiples of 32, heaviest class first
  cstart[0] = 0;
  #pragma unroll
  for (int c = 0; c < RT_NCLS; c++) cstart[c + 1] = cstart[c] + ((qin.ctl[RT_NCLS - c] + 31) & ~31);
  // The loop must be provably warp-uniform for ptxas (see the note at the stand-in lanes below).  Two things it needs:
  // a trip count bounded by a kernel parameter (with `for (;;)` it proves nothing and guards every warp collective of
  // the kernel with a divergence check, BRA.DIV in the SASS; the bound itself is never reached, a warp cannot claim
  // more blocks than the queue has entries), and an exit test on lane-0 broadcasts only.
  const int vend = __shfl_sync(0xffffffffu, cstart[RT_NCLS], 0);
  for (int claim_no = 0; claim_no <= qin.cap; claim_no++) {
    int vbase = 0;
    if ((threadIdx.x & 31) == 0) vbase = atomicAdd(&qin.ctl[RT_CTL_CURSOR], 32);
    vbase = __shfl_sync(0xffffffffu, vbase, 0);
    if (vbase >= vend) break;
    int c = 0;
    #pragma unroll
    for (int k = 1; k < RT_NCLS; k++) c += (vbase >= cstart[k]) ? 1 : 0;
    const int cls = RT_NCLS - 1 - c;
    // Lanes past the end of the class list do NOT leave the loop body: they shade a copy of lane 0's entry with every
    // ray switched off and write nothing.  The body then runs in warp-uniform control flow, which is what lets ptxas
    // keep the scan's sphere operands in uniform registers (with a divergent `continue` here it used none: every
    // FFMA of the scan read three register-file operands and issued at 0.61 per clock).
    int k_in = vbase - cstart[c] + (threadIdx.x & 31);
    const bool have = k_in < qin.ctl[1 + cls];
    constexpr unsigned vmask = 0xffffffffu;
    if (!have) k_in = vbase - cstart[c];                      // lane 0's entry: a padded block holds at least one
    const int item = qin.cls[(size_t)cls * qin.cap + k_in];
    const float4 r0 = qin.r0[item], r1 = qin.r1[item], r2 = qin.r2[item];
    const v3 P = mk3(r0.x, r0.y, r0.z), N = mk3(r1.x, r1.y, r1.z), V = mk3(r2.x, r2.y, r2.z);
    const int pix = __float_as_int(r0.w), obj = __float_as_int(r1.w);
    const float tr = r2.w;
    const float4 m0 = sc.mat[2 * obj], m1 = sc.mat[2 * obj + 1];
    const float refl = m0.w, opacity = m1.w;
    int obj_sph = -1;                                          // sphere slot of obj (accel mode skips by slot)
    if constexpr (ACCEL) { const int2 osl = sc.slot[obj]; obj_sph = osl.x == 0 ? osl.y : -1; }
    const bool need_refl = have && (refl > 0.0f) && (level < sc.max_depth - 1);   // C/raycast.c:349,352
    const bool lit = have && opacity > 0.0f;                 // C/raycast.c:385
    if (have) n_hp++;
    if (need_refl) n_refl++;
    if (lit) n_lit++;

    const float p2s = RT_P2_EPS * __fmaf_rn(P.z, P.z, __fmaf_rn(P.y, P.y, P.x * P.x));   // margin term of scan_chunk_c
    v3 rd = mk3(0.f, 0.f, 0.f);
    if (need_refl) rd = norm3(reflect3(V, N));               // C/raycast.c:353-355
    float best_t = RT_INF;
    int best_idx = -1;
    v3 out = mk3(0.f, 0.f, 0.f);

    const int n_pass = (sc.n_lights + LB - 1) / LB > 0 ? (sc.n_lights + LB - 1) / LB : 1;
    for (int pass = 0; pass < n_pass; pass++) {
      // ---- ray setup ---------------------------------------------------------------
      // Slots are packed: the reflection ray (when there is one) takes slot 0, then one slot
      // per light whose shadow ray can matter.  A light behind the surface (N.L <= 0) gets
      // no slot: diffuse_light and specular_light both return black for it
      // (C/raycast.c:714-717, 739-742), so whatever its shadow ray finds, the light adds
      // exactly +0 to the colour (its attenuations are finite: the host clears the cull flag
      // of any light for which that is not certain).  The ray still counts as a ray of the
      // reference semantics (it is cast there, :401); only its work is skipped, like the
      // rest of a shadow search after its first hit.
      // Directions go to shared memory (dsm): the sphere scan loads them into registers,
      // every other section indexes them at run time.
      const bool r_on = need_refl && pass == 0;
      unsigned alive = 0;
      unsigned long long lmap = 0;                           // slot -> light of this pass, 4 bits each
      int cnt = 0;
      if (r_on) { dsm[0][tid] = rd.x; dsm[1][tid] = rd.y; dsm[2][tid] = rd.z; alive = 1u; cnt = 1; }
      if (lit) {
        #pragma unroll 1
        for (int l = pass * LB; l < min(pass * LB + LB, sc.n_lights); l++) {
          const float4 lp = sc.lights[4 * l];
          const v3 dj = norm3(sub3(mk3(lp.x, lp.y, lp.z), P));   // C/raycast.c:396-398
          if (sc.lights[4 * l + 3].w != 0.0f && dot3(N, dj) <= 0.0f) continue;
          dsm[3 * cnt][tid] = dj.x; dsm[3 * cnt + 1][tid] = dj.y; dsm[3 * cnt + 2][tid] = dj.z;
          alive |= 1u << cnt;
          lmap |= (unsigned long long)(l - pass * LB) << (4 * (cnt - (r_on ? 1 : 0)));   // 16 shadow slots x 4 bits
          cnt++;
        }
      }
      const int wmax = __reduce_max_sync(vmask, cnt);
      int nrv = pick_variant(wmax);
      #pragma unroll 1
      // unused slots of the chosen scan variant hold NaN: they lose every max and pass no test
      for (int j = cnt; j < nrv; j++) { dsm[3 * j][tid] = RT_NAN; dsm[3 * j + 1][tid] = RT_NAN; dsm[3 * j + 2][tid] = RT_NAN; }
      if (have) n_pass_sum++;   // stand-in lanes execute the pass but their tests are not counted as work
      // A shadow ray is done at its first hit (find_nearest_object's result is only compared with
      // NULL, C/raycast.c:401-402) and the order of the tests is free, so: planes first (few, and
      // every one cuts half of space), spheres by descending radius (host order), and whenever
      // the largest number of rays still alive in the warp fits a smaller scan variant the
      // survivors are packed to the front of the slot array and the scan continues with it.
      auto repack = [&]() {
        const int wlive = __reduce_max_sync(vmask, __popc(alive));
        const int nv = pick_variant(wlive);
        if (nv >= nrv) return;                                 // warp-uniform
        repack_slots(dsm, tid, r_on ? 1 : 0, nv, alive, lmap);
        nrv = nv;
      };
      // ---- planes ----------------------------------------------------------------
      // t = (-1*num)/den as the reference computes it (C/raycast.c:545-562): den == 0 or
      // t < 0 is a miss, and a hit counts only with 0 < t < inf (:476).
      d_pln += (unsigned)__popc(alive) * (unsigned)sc.n_planes;
      for (int p = 0; p < sc.n_planes; p++) {
        const int idx = sc.pln_idx[p];
        if (idx == obj) continue;                              // C/raycast.c:450
        const float4 pp = sc.pln[2 * p], pn = sc.pln[2 * p + 1];
        const v3 n = mk3(pn.x, pn.y, pn.z);
        const float nn = mul32(-1.0f, dot3(sub3(P, mk3(pp.x, pp.y, pp.z)), n));
        for (unsigned m = alive; m; m &= m - 1) {
          const int j = __ffs(m) - 1;
          const v3 dj = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
          const float den = dot3(dj, n);
          if (r_on && j == 0) {
            const float t = __fdiv_rn(nn, den);
            if (den != 0.0f && better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; }
          } else {
            // A shadow ray only asks whether 0 < t < inf.  The sign of an IEEE quotient is
            // exact, so with both magnitudes in [2^-40, 2^40] (no underflow to 0, no overflow
            // to inf) that is "nn and den have the same sign"; otherwise divide.
            const float an = fabsf(nn), ad = fabsf(den);
            bool hit;
            if (an >= 9.094947017729282e-13f && an <= 1.099511627776e12f && ad >= 9.094947017729282e-13f && ad <= 1.099511627776e12f)
              hit = (nn > 0.0f) == (den > 0.0f);
            else {
              const float t = __fdiv_rn(nn, den);
              hit = den != 0.0f && t > 0.0f && t < RT_INF;
            }
            if (hit) alive &= ~(1u << j);
          }
        }
      }
      if (sc.n_planes) repack();
      // ---- spheres ---------------------------------------------------------------
      if constexpr (!ACCEL) {
        if (have) o_sph += (unsigned)sc.n_spheres;
        for (int cs = 0; cs < sc.n_spheres; cs += RT_CHUNK) {
          const int n = min(RT_CHUNK, sc.n_spheres - cs);
          d_sph += (unsigned)__popc(alive) * (unsigned)n;
          if (have) d_slots += (unsigned)nrv * (unsigned)n;
          if (cs + n <= sc.n_csph) scan_spheres_v<NR, VMIN>(nrv, dsm, tid, P, p2s, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand);
          else scan_spheres_g<NR, VMIN>(nrv, dsm, tid, P, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand);
          if (nrv > VMIN) repack();
        }
      } else {
        // shadow rays: only the spheres listed in the cell of the light buffer that the direction
        // from the light to P falls in (level by |P - L|; beyond the last level: every sphere)
        auto shadow_query = [&](const v3 Pq, const v3 dj, const int l, const int osph) -> bool {
          const float4 lp = sc.lights[4 * l];
          const float wx = lp.x - Pq.x, wy = lp.y - Pq.y, wz = lp.z - Pq.z;
          const float q = __fmaf_rn(wz, wz, __fmaf_rn(wy, wy, wx * wx));
          int lv = -1;
          #pragma unroll
          for (int v = RT_LB_LEVELS - 1; v >= 0; v--) lv = q <= sc.lb_caps[l * RT_LB_LEVELS + v] ? v : lv;
          const int cell = lb_cell(-dj.x, -dj.y, -dj.z);
          bool occl = false;
          if (lv >= 0 && cell >= 0) {
            const unsigned *st = sc.lb_start + ((size_t)(l * RT_LB_LEVELS + lv) * 6 * RT_LB_RES * RT_LB_RES + cell);
            const unsigned k1 = st[1];
            for (unsigned k = st[0]; k < k1 && !occl; k++) {
              const int ss = sc.lb_items[k];
              const float4 hh = sc.lb_hot[k];                   // the sphere's filter operands, stored with the list
              o_sph++; d_sph++;
              if (ss != osph) occl = shadow_hits_sphere(sc, Pq, dj, ss, hh, n_cand);   // C/raycast.c:450
            }
          } else {
            unsigned nt = 0;
            float bt = RT_INF;
            int bi = -1;
            occl = bvh_trace<true>(sc, Pq, dj, osph, bt, bi, n_cand, nt);
            o_sph += nt; d_sph += nt;
          }
          return occl;
        };
        // (dealing the warp's (hit point, ray) pairs out over its lanes by shuffle -- to even out the lanes' different
        // numbers of live shadow rays -- was tried and dropped: 425 instead of 279 ms on C4, the owner look-up and the
        // loss of per-lane locality cost more than the idle lanes)
        for (unsigned m = alive & ~(r_on ? 1u : 0u); m; m &= m - 1) {
          const int j = __ffs(m) - 1;
          const int l = pass * LB + (int)((lmap >> (4 * (j - (r_on ? 1 : 0)))) & 15ull);
          const v3 dj = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
          if (shadow_query(P, dj, l, obj_sph)) alive &= ~(1u << j);
        }
        // reflection ray: nearest hit through the bounding-sphere hierarchy
        if (r_on) {
          unsigned nt = 0;
          bvh_trace<false>(sc, P, rd, obj_sph, best_t, best_idx, n_cand, nt);
          o_sph += nt; d_sph += nt; d_slots += nt;
        }
        repack();
      }
      // ---- quadrics --------------------------------------------------------------
      if constexpr (QUADS) {
        int q0 = 0;
        if constexpr (ACCEL) {
          // indexed ellipsoids through their hierarchy (rt_qindex.cuh); every other quadric is scanned below
          if (sc.n_qidx > 0) {
            q0 = sc.n_qidx;
            for (unsigned m = alive & ~(r_on ? 1u : 0u); m; m &= m - 1) {
              const int j = __ffs(m) - 1;
              const v3 dj = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
              unsigned nt = 0;
              bool occl = false;
              int lv = -1, cell = -1;
              if (sc.lbq_start != nullptr) {
                // the ellipsoids listed in the cell of the light's cube map that the direction from the light to P
                // falls in (level by |P - L|), as for the spheres
                const int l = pass * LB + (int)((lmap >> (4 * (j - (r_on ? 1 : 0)))) & 15ull);
                const float4 lp = sc.lights[4 * l];
                const float wx = lp.x - P.x, wy = lp.y - P.y, wz = lp.z - P.z;
                const float q = __fmaf_rn(wz, wz, __fmaf_rn(wy, wy, wx * wx));
                #pragma unroll
                for (int v = RT_LB_LEVELS - 1; v >= 0; v--) lv = q <= sc.lbq_caps[l * RT_LB_LEVELS + v] ? v : lv;
                cell = lb_cell(-dj.x, -dj.y, -dj.z);
                if (lv >= 0 && cell >= 0) {
                  const unsigned *st = sc.lbq_start + ((size_t)(l * RT_LB_LEVELS + lv) * 6 * RT_LB_RES * RT_LB_RES + cell);
                  const unsigned k1 = st[1];
                  for (unsigned k = st[0]; k < k1 && !occl; k++) {
                    const int kk = sc.lbq_items[k];
                    nt++;
                    float t;
                    if (sc.quad_idx[kk] != obj && quadric_reported<true>(sc, P, dj, kk, t, n_cand)) occl = t > 0.0f && t < RT_INF;   // C/raycast.c:450
                  }
                }
              }
              if (!(lv >= 0 && cell >= 0)) {
                float bt = RT_INF;
                int bi = -1;
                occl = qbvh_trace<true, true>(sc, P, dj, obj, bt, bi, n_cand, nt);
              }
              if (occl) alive &= ~(1u << j);
              d_quad += nt; o_quad += nt;
            }
            if (r_on) {
              unsigned nt = 0;
              qbvh_trace<false, true>(sc, P, rd, obj, best_t, best_idx, n_cand, nt);
              d_quad += nt; o_quad += nt; d_slots += nt;
            }
            repack();
          }
        }
        for (int cs = q0; cs < sc.n_quadrics; cs += RT_CHUNK) {
          const int n = min(RT_CHUNK, sc.n_quadrics - cs);
          // counted as executed: the bounding-ball tests (sphere-type tests: d_ball / o_ball) and, for the survivors, the
          // quadric filter set-ups and per-ray evaluations (o_quad / d_quad, counted inside resolve_quadrics)
          d_ball += (unsigned)__popc(alive) * (unsigned)n;
          if (have) { d_slots += (unsigned)nrv * (unsigned)n; o_ball += (unsigned)n; }
          scan_quadrics_v<NR, VMIN>(nrv, dsm, tid, P, sc, cs, n, obj, r_on, alive, best_t, best_idx, n_cand, o_quad, d_quad);
          if (nrv > VMIN) repack();
        }
      }
      // ---- Phong for the lights of this pass that stayed unoccluded, in light order
      for (unsigned m = alive & ~(r_on ? 1u : 0u); m; m &= m - 1) {
        const int j = __ffs(m) - 1;
        n_ls++;
        const int l = pass * LB + (int)((lmap >> (4 * (j - (r_on ? 1 : 0)))) & 15ull);
        const float4 l0 = sc.lights[4 * l], l1 = sc.lights[4 * l + 1], l2 = sc.lights[4 * l + 2];
        const v3 Lpos = mk3(l0.x, l0.y, l0.z);
        const v3 Ld = mk3(dsm[3 * j][tid], dsm[3 * j + 1][tid], dsm[3 * j + 2][tid]);
        const float dist = len3(sub3(Lpos, P));              // C/raycast.c:397
        // radial_attenuation, C/raycast.c:666-669
        const float rad = (float)__ddiv_rn(1.0, add64((double)add32(l2.x, mul32(l2.y, dist)),
                                                     mul64((double)l2.z, mul64((double)dist, (double)dist))));
        // angular_attenuation, C/raycast.c:679-696
        float ang = 1.0f;
        if (__float_as_int(l0.w) == 1) {
          const float4 l3 = sc.lights[4 * l + 3];
          const v3 vo = norm3(sub3(P, Lpos));
          const float alpha = dot3(vo, mk3(l3.x, l3.y, l3.z));
          ang = (alpha < l1.w) ? 0.0f : (float)pow_a0((double)alpha, l2.w);
        }
        // diffuse_light :708-720 and specular_light :733-758
        v3 diff = mk3(0.f, 0.f, 0.f), spec = mk3(0.f, 0.f, 0.f);
        const float theta = dot3(N, Ld);
        if (!(theta <= 0.0f)) {
          diff = mk3(mul32(mul32(m0.x, l1.x), theta), mul32(mul32(m0.y, l1.y), theta), mul32(mul32(m0.z, l1.z), theta));
          const v3 R = reflect3(Ld, N);
          const v3 nV = mk3(mul32(V.x, -1.0f), mul32(V.y, -1.0f), mul32(V.z, -1.0f));
          const double angle = (double)dot3(nV, R);
          if (!(angle > 0.0)) {
            const double p = pow20(angle);
            spec = mk3((float)mul64((double)mul32(m1.x, l1.x), p), (float)mul64((double)mul32(m1.y, l1.y), p),
                       (float)mul64((double)mul32(m1.z, l1.z), p));
          }
        }
        out.x = add32(out.x, mul32(mul32(add32(diff.x, spec.x), rad), ang));   // C/raycast.c:414-416
        out.y = add32(out.y, mul32(mul32(add32(diff.y, spec.y), rad), ang));
        out.z = add32(out.z, mul32(mul32(add32(diff.z, spec.z), rad), ang));
      }
    }
    if (lit) out = mk3(mul32(out.x, opacity), mul32(out.y, opacity), mul32(out.z, opacity)); // :419
    if (!have) {
      // nothing to write for a stand-in lane
    } else if (lvl != nullptr) {
      // RT_MODE_SHOOT: colour and reflectivity of this hit, combined from the deepest level outwards by k_finalize
      // (C/raycast.c:291-294); lvl = this level's slice of the per-tile buffer
      lvl[pix] = make_float4(out.x, out.y, out.z, refl);
    } else if (level == 0) {
      c0[pix] = make_float4(out.x, out.y, out.z, 0.f);        // added last, C/raycast.c:377-378
    } else {
      float4 ac = acc[pix];                                    // C/raycast.c:361-363
      ac.x = add32(ac.x, mul32(out.x, tr)); ac.y = add32(ac.y, mul32(out.y, tr)); ac.z = add32(ac.z, mul32(out.z, tr));
      acc[pix] = ac;
    }
    {
      const bool bounce = need_refl && best_idx >= 0;
      v3 P2 = P, N2 = N;
      int cls2 = 0;
      if (bounce) {
        finish_hit(sc, P, rd, best_t, best_idx, P2, N2);
        cls2 = hit_class(sc, P2, N2, best_idx, level + 1 < sc.max_depth - 1, NR);
      }
      queue_push(qout, bounce, make_float4(P2.x, P2.y, P2.z, __int_as_float(pix)), make_float4(N2.x, N2.y, N2.z, __int_as_float(best_idx)),
                 make_float4(rd.x, rd.y, rd.z, mul32(tr, refl)), cls2);                  // total_refl, :345,362
    }
  }
  unsigned v[14] = {n_refl, n_lit, n_hp, n_pass_sum, d_sph, n_ls, n_cand, d_pln, d_quad, d_slots, o_sph, o_quad, d_ball, o_ball};
  #pragma unroll
  for (int k = 0; k < 14; k++) {
    #pragma unroll
    for (int off = 16; off; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
  }
  if ((threadIdx.x & 31) == 0) {
    typedef unsigned long long u64;
    const u64 np = sc.n_planes;
    if (v[0]) atomicAdd(&counters[CT_RAYS_REFLECTION], (u64)v[0]);
    if (v[1]) atomicAdd(&counters[CT_RAYS_SHADOW], (u64)v[1] * sc.n_lights);
    if (v[2]) atomicAdd(&counters[CT_HIT_POINTS], (u64)v[2]);
    if (v[3] && np) atomicAdd(&counters[CT_PLN_O], v[3] * np);
    if (v[11]) atomicAdd(&counters[CT_QUAD_O], (u64)v[11]);
    if (v[4]) atomicAdd(&counters[CT_SPH_D], (u64)v[4]);
    if (v[7]) atomicAdd(&counters[CT_PLN_D], (u64)v[7]);
    if (v[8]) atomicAdd(&counters[CT_QUAD_D], (u64)v[8]);
    if (v[9]) atomicAdd(&counters[CT_SLOTS], (u64)v[9]);
    if (v[10]) atomicAdd(&counters[CT_SPH_O], (u64)v[10]);
    if (v[5]) atomicAdd(&counters[CT_LIGHT_SAMPLES], (u64)v[5]);
    if (v[6]) atomicAdd(&counters[CT_CANDIDATES], (u64)v[6]);
    if (v[12]) atomicAdd(&counters[CT_BALL_D], (u64)v[12]);
    if (v[13]) atomicAdd(&counters[CT_BALL_O], (u64)v[13]);
  }
}

// ---------------------------------------------------------------------------------
// k_finalize: out = acc + primary (C/raycast.c:377-378), *255, ppm_clamp, truncating
// float->uint8 (C/raycast.c:122-126; NaN -> 0 like x86 cvttss2si's low byte).
// One thread per 4 consecutive pixels of the rank buffer = 12 bytes = 3 aligned 32-bit
// stores; a warp writes 384 contiguous bytes.
// ---------------------------------------------------------------------------------
RT_DEV unsigned quant8(float v) {
  v = mul32(v, 255.0f);
  if (v > 255.0f) v = 255.0f;
  if (v < 0.0f) v = 0.0f;
  return __float2uint_rz(v) & 0xffu;   // NaN -> 0
}

// RT_MODE_SHOOT (lvl != NULL): the pixel is the nested sum of shoot(), C/raycast.c:291-294,
//   out_k = colour_k + out_(k+1) * reflectivity_k,  out beyond the last hit = black,
// over the per-level records k_shade left in lvl[k * lvl_stride + pixel] (zero where the chain has no hit).
__global__ void __launch_bounds__(256)
k_finalize(TileGeom tg, const float4 *__restrict__ acc, const float4 *__restrict__ c0,
           uint8_t *__restrict__ rank_buf, const float4 *__restrict__ lvl, size_t lvl_stride, int depth) {
  // groups are aligned on the rank buffer (pixel index multiple of 4 => byte offset multiple of 12)
  const long long g_first = tg.first_px >> 2, g_last = (tg.first_px + tg.n_px - 1) >> 2;
  const long long n_groups = g_last - g_first + 1;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x; gi < n_groups; gi += stride) {
    const long long px0 = (g_first + gi) << 2;
    unsigned b[12];
    bool full = true;
    #pragma unroll
    for (int k = 0; k < 4; k++) {
      long long i = px0 + k - tg.first_px;
      if (i >= 0 && i < tg.n_px) {
        if (lvl != nullptr) {
          float ox = 0.0f, oy = 0.0f, oz = 0.0f;
          for (int l = depth - 1; l >= 0; l--) {
            const float4 r = lvl[(size_t)l * lvl_stride + i];
            ox = add32(r.x, mul32(ox, r.w)); oy = add32(r.y, mul32(oy, r.w)); oz = add32(r.z, mul32(oz, r.w));
          }
          b[3 * k] = quant8(ox); b[3 * k + 1] = quant8(oy); b[3 * k + 2] = quant8(oz);
        } else {
          float4 a = acc[i], c = c0[i];
          b[3 * k] = quant8(add32(a.x, c.x)); b[3 * k + 1] = quant8(add32(a.y, c.y)); b[3 * k + 2] = quant8(add32(a.z, c.z));
        }
      } else { full = false; b[3 * k] = b[3 * k + 1] = b[3 * k + 2] = 0; }
    }
    uint8_t *dst = rank_buf + px0 * 3;
    if (full) {
      uint32_t *d32 = reinterpret_cast<uint32_t *>(dst);
      d32[0] = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
      d32[1] = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
      d32[2] = b[8] | (b[9] << 8) | (b[10] << 16) | (b[11] << 24);
    } else {
      #pragma unroll
      for (int k = 0; k < 4; k++) {
        long long i = px0 + k - tg.first_px;
        if (i >= 0 && i < tg.n_px) { dst[3 * k] = (uint8_t)b[3 * k]; dst[3 * k + 1] = (uint8_t)b[3 * k + 1]; dst[3 * k + 2] = (uint8_t)b[3 * k + 2]; }
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// k_assemble: de-interleave the gathered rank buffers (rank r's rows back to back at
// staging + rank_off[r]) into the row-major frame.  16-byte copies when the row pitch
// allows it, bytes otherwise.
// ---------------------------------------------------------------------------------
struct AssembleArgs {
  long long rank_off[8];
  int width, height, nranks, band_h;
};

__global__ void __launch_bounds__(256)
k_assemble(AssembleArgs a, const uint8_t *__restrict__ staging, uint8_t *__restrict__ frame) {
  const long long pitch = (long long)a.width * 3;
  const bool vec = (pitch % 16) == 0;
  const long long per_row = vec ? pitch / 16 : pitch;
  const long long total = per_row * a.height;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int y = (int)(i / per_row);
    const long long e = i - (long long)y * per_row;
    const int band = y / a.band_h, r = band % a.nranks;
    const int lr = (band / a.nranks) * a.band_h + (y - band * a.band_h);
    const uint8_t *src = staging + a.rank_off[r] + (long long)lr * pitch;
    uint8_t *dst = frame + (long long)y * pitch;
    if (vec) reinterpret_cast<uint4 *>(dst)[e] = reinterpret_cast<const uint4 *>(src)[e];
    else dst[e] = src[e];
  }
}
