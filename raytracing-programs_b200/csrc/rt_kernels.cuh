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

#define RT_INF __int_as_float(0x7f800000)

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

RT_DEV int queue_push_slot(int *count) {
  unsigned m = __activemask();
  int lane = threadIdx.x & 31;
  int leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(count, __popc(m));
  base = __shfl_sync(m, base, leader);
  return base + __popc(m & ((1u << lane) - 1u));
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
__global__ void __launch_bounds__(256)
k_primary(DevScene sc, TileGeom tg, float4 *__restrict__ acc, float4 *__restrict__ c0,
          HitQueue q, unsigned long long *__restrict__ counters) {
  const float pw = __fdiv_rn(sc.cam_w, (float)tg.width);    // C/raycast.c:110
  const float ph = __fdiv_rn(sc.cam_h, (float)tg.height);   // C/raycast.c:109
  const double half_w = __ddiv_rn((double)sc.cam_w, 2.0), half_h = __ddiv_rn((double)sc.cam_h, 2.0);
  unsigned long long n_hits = 0, n_sph = 0, n_pln = 0, n_quad = 0, n_cand = 0;
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < tg.n_px; i += stride) {
    long long g = tg.first_px + i;
    int lr = (int)(g / tg.width), x = (int)(g - (long long)lr * tg.width);
    int y = local_to_image_row(lr, tg.rank, tg.nranks, tg.band_h);
    // C/raycast.c:115-118 (double expression, float store, normalise)
    v3 d;
    d.x = (float)add64(sub64(0.0, half_w), mul64((double)pw, add64((double)x, 0.5)));
    d.y = (float)sub64(add64(0.0, half_h), mul64((double)ph, add64((double)y, 0.5)));
    d.z = -1.0f;
    d = norm3(d);
    const v3 o = mk3(0.0f, 0.0f, 0.0f);
    const float a = ray_a(d);
    float best_t = RT_INF;
    int best_idx = -1;
    // spheres: origin part precomputed on the host (origin is exactly 0)
    for (int s = 0; s < sc.n_spheres; s++) {
      float4 sp = sc.sph_prim[s];
      float dt = dot3(d, mk3(sp.x, sp.y, sp.z));
      float dq = __fmaf_rn(dt, dt, -mul32(a, sp.w));
      if (dq >= 0.0f) {
        n_cand++;
        float t;
        if (sphere_t_literal(o, d, sc.sph[s], t)) {
          int idx = sc.sph_idx[s];
          if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; }
        }
      }
    }
    n_sph += sc.n_spheres;
    for (int p = 0; p < sc.n_planes; p++) {
      float4 pp = sc.pln[2 * p], pn = sc.pln[2 * p + 1];
      v3 n = mk3(pn.x, pn.y, pn.z);
      float num = dot3(sub3(o, mk3(pp.x, pp.y, pp.z)), n);
      float t;
      if (plane_t(num, d, n, t)) {
        int idx = sc.pln_idx[p];
        if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; }
      }
    }
    n_pln += sc.n_planes;
    for (int k = 0; k < sc.n_quadrics; k++) {
      quad10 qq = load_quad(sc.quad, k);
      float c_q = (float)quadric_cq(o, qq);
      float t;
      if (quadric_t(o, d, qq, c_q, t)) { // skip_index == -1: no z rule on primary rays
        int idx = sc.quad_idx[k];
        if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; }
      }
    }
    n_quad += sc.n_quadrics;
    acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (best_idx < 0) {
      c0[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      n_hits++;
      v3 P, N;
      finish_hit(sc, o, d, best_t, best_idx, P, N);
      int slot = queue_push_slot(q.count);
      q.r0[slot] = make_float4(P.x, P.y, P.z, __int_as_float(i));
      q.r1[slot] = make_float4(N.x, N.y, N.z, __int_as_float(best_idx));
      q.r2[slot] = make_float4(d.x, d.y, d.z, 1.0f);
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
  }
  if ((threadIdx.x & 31) == 0) {
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
template <int LB>
__global__ void __launch_bounds__(128)
k_shade(DevScene sc, int level, HitQueue qin, HitQueue qout, float4 *__restrict__ acc,
        float4 *__restrict__ c0, unsigned long long *__restrict__ counters) {
  constexpr int NR = LB + 1;
  unsigned long long n_refl = 0, n_shadow = 0, n_hp = 0, n_so = 0, n_sd = 0, n_po = 0, n_pd = 0,
                     n_qo = 0, n_qd = 0, n_ls = 0, n_cand = 0;
  const int n_items = *qin.count;
  const int stride = gridDim.x * blockDim.x;
  for (int item = blockIdx.x * blockDim.x + threadIdx.x; item < n_items; item += stride) {
    const float4 r0 = qin.r0[item], r1 = qin.r1[item], r2 = qin.r2[item];
    const v3 P = mk3(r0.x, r0.y, r0.z), N = mk3(r1.x, r1.y, r1.z), V = mk3(r2.x, r2.y, r2.z);
    const int pix = __float_as_int(r0.w), obj = __float_as_int(r1.w);
    const float tr = r2.w;
    const float4 m0 = sc.mat[2 * obj], m1 = sc.mat[2 * obj + 1];
    const float refl = m0.w, opacity = m1.w;
    const bool need_refl = (refl > 0.0f) && (level < 6);   // C/raycast.c:349,352
    const bool lit = opacity > 0.0f;                         // C/raycast.c:385
    n_hp++;
    if (need_refl) n_refl++;
    if (lit) n_shadow += sc.n_lights;

    v3 rd = mk3(0.f, 0.f, 0.f);
    if (need_refl) rd = norm3(reflect3(V, N));               // C/raycast.c:353-355
    float best_t = RT_INF;
    int best_idx = -1;
    v3 out = mk3(0.f, 0.f, 0.f);

    const int n_pass = (sc.n_lights + LB - 1) / LB > 0 ? (sc.n_lights + LB - 1) / LB : 1;
    for (int pass = 0; pass < n_pass; pass++) {
      // ---- ray setup -------------------------------------------------------------
      v3 d[NR];
      float a[NR];
      unsigned alive = 0;
      d[0] = mk3(0.f, 0.f, 0.f); a[0] = 1.0f;
      if (need_refl && pass == 0) { d[0] = rd; a[0] = ray_a(rd); alive |= 1u; }
      #pragma unroll
      for (int j = 1; j < NR; j++) {
        int l = pass * LB + j - 1;
        d[j] = mk3(0.f, 0.f, 0.f); a[j] = 1.0f;
        if (lit && l < sc.n_lights) {
          float4 lp = sc.lights[4 * l];
          d[j] = norm3(sub3(mk3(lp.x, lp.y, lp.z), P));     // C/raycast.c:396-398
          a[j] = ray_a(d[j]);
          alive |= 1u << j;
        }
      }
      const int n_live = __popc(alive);
      // ---- spheres ---------------------------------------------------------------
      #pragma unroll 2
      for (int s = 0; s < sc.n_spheres; s++) {
        const float4 sp = sc.sph[s];
        const v3 tmp = sub3(P, mk3(sp.x, sp.y, sp.z));
        const float c = __fmaf_rn(-sp.w, sp.w, dot3(tmp, tmp));
        float m = -1.0f;
        #pragma unroll
        for (int j = 0; j < NR; j++) {
          float dt = dot3(d[j], tmp);
          float dq = __fmaf_rn(dt, dt, -mul32(a[j], c));
          m = fmaxf(m, dq);
        }
        if (m >= 0.0f) {
          const int idx = sc.sph_idx[s];
          if (idx != obj) {                                  // C/raycast.c:450
            #pragma unroll
            for (int j = 0; j < NR; j++) {
              float dt = dot3(d[j], tmp);
              float dq = __fmaf_rn(dt, dt, -mul32(a[j], c));
              if (dq >= 0.0f && ((alive >> j) & 1u)) {
                n_cand++;
                float t;
                if (sphere_t_literal(P, d[j], sp, t)) {
                  if (j == 0) { if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; } }
                  else if (t > 0.0f && t < RT_INF) alive &= ~(1u << j);
                }
              }
            }
          }
        }
      }
      n_so += sc.n_spheres; n_sd += (unsigned long long)sc.n_spheres * n_live;
      // ---- planes ----------------------------------------------------------------
      for (int p = 0; p < sc.n_planes; p++) {
        const int idx = sc.pln_idx[p];
        if (idx == obj) continue;
        const float4 pp = sc.pln[2 * p], pn = sc.pln[2 * p + 1];
        const v3 n = mk3(pn.x, pn.y, pn.z);
        const float num = dot3(sub3(P, mk3(pp.x, pp.y, pp.z)), n);
        #pragma unroll
        for (int j = 0; j < NR; j++) {
          float t;
          if (((alive >> j) & 1u) && plane_t(num, d[j], n, t)) {
            if (j == 0) { if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; } }
            else if (t > 0.0f && t < RT_INF) alive &= ~(1u << j);
          }
        }
      }
      n_po += sc.n_planes; n_pd += (unsigned long long)sc.n_planes * n_live;
      // ---- quadrics --------------------------------------------------------------
      for (int k = 0; k < sc.n_quadrics; k++) {
        const int idx = sc.quad_idx[k];
        if (idx == obj) continue;
        const quad10 qq = load_quad(sc.quad, k);
        const float c_q = (float)quadric_cq(P, qq);
        #pragma unroll
        for (int j = 0; j < NR; j++) {
          float t;
          if (((alive >> j) & 1u) && quadric_t(P, d[j], qq, c_q, t)) {
            if (add32(P.z, mul32(t, d[j].z)) < P.z) continue;   // C/raycast.c:492-494 (skip_index != -1 here)
            if (j == 0) { if (better_hit(t, idx, best_t, best_idx)) { best_t = t; best_idx = idx; } }
            else if (t > 0.0f && t < RT_INF) alive &= ~(1u << j);
          }
        }
      }
      n_qo += sc.n_quadrics; n_qd += (unsigned long long)sc.n_quadrics * n_live;
      // ---- Phong for the lights of this pass that stayed unoccluded, in light order
      #pragma unroll
      for (int j = 1; j < NR; j++) {
        if (!((alive >> j) & 1u)) continue;
        n_ls++;
        const int l = pass * LB + j - 1;
        const float4 l0 = sc.lights[4 * l], l1 = sc.lights[4 * l + 1], l2 = sc.lights[4 * l + 2];
        const v3 Lpos = mk3(l0.x, l0.y, l0.z);
        const v3 Ld = d[j];
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
    if (level == 0) {
      c0[pix] = make_float4(out.x, out.y, out.z, 0.f);        // added last, C/raycast.c:377-378
    } else {
      float4 ac = acc[pix];                                    // C/raycast.c:361-363
      ac.x = add32(ac.x, mul32(out.x, tr)); ac.y = add32(ac.y, mul32(out.y, tr)); ac.z = add32(ac.z, mul32(out.z, tr));
      acc[pix] = ac;
    }
    if (need_refl && best_idx >= 0) {
      v3 P2, N2;
      finish_hit(sc, P, rd, best_t, best_idx, P2, N2);
      const int slot = queue_push_slot(qout.count);
      qout.r0[slot] = make_float4(P2.x, P2.y, P2.z, __int_as_float(pix));
      qout.r1[slot] = make_float4(N2.x, N2.y, N2.z, __int_as_float(best_idx));
      qout.r2[slot] = make_float4(rd.x, rd.y, rd.z, mul32(tr, refl));   // total_refl, :345,362
    }
  }
  unsigned long long v[11] = {n_refl, n_shadow, n_hp, n_so, n_sd, n_po, n_pd, n_qo, n_qd, n_ls, n_cand};
  const int where[11] = {CT_RAYS_REFLECTION, CT_RAYS_SHADOW, CT_HIT_POINTS, CT_SPH_O, CT_SPH_D, CT_PLN_O,
                         CT_PLN_D, CT_QUAD_O, CT_QUAD_D, CT_LIGHT_SAMPLES, CT_CANDIDATES};
  #pragma unroll
  for (int k = 0; k < 11; k++) {
    #pragma unroll
    for (int off = 16; off; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
    if ((threadIdx.x & 31) == 0 && v[k]) atomicAdd(&counters[where[k]], v[k]);
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

__global__ void __launch_bounds__(256)
k_finalize(TileGeom tg, const float4 *__restrict__ acc, const float4 *__restrict__ c0,
           uint8_t *__restrict__ rank_buf) {
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
        float4 a = acc[i], c = c0[i];
        b[3 * k] = quant8(add32(a.x, c.x)); b[3 * k + 1] = quant8(add32(a.y, c.y)); b[3 * k + 2] = quant8(add32(a.z, c.z));
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
