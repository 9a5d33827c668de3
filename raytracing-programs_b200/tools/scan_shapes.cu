// tools/scan_shapes.cu -- microbenchmark of candidate inner-loop shapes for the sphere reject scan of k_shade
// (DESIGN.md "sphere reject filter").  Not product code: it decides which loop structure the kernel gets.
//   A  thread per hit point, NRV rays in registers, spheres streamed with warp-uniform loads (round-1 structure)
//   B  warp per hit point, lanes over spheres (S per lane), rays streamed from shared memory by broadcast loads,
//      two rays per trip (one 3-input max per sphere and trip); the ray count is exact per hit point
//   C  B with the FMA-pipe work packed over sphere pairs (mul/fma.rn.f32x2, sm_100+)
// Reports tests/s and the implied FMA-pipe utilisation (origin part 9 + 3 per ray and sphere, scalar-equivalent).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define EPS 3.814697265625e-06f

__device__ __forceinline__ u64 pk(float lo, float hi) { return ((u64)__float_as_uint(hi) << 32) | __float_as_uint(lo); }
__device__ __forceinline__ float lo32(u64 v) { return __uint_as_float((unsigned)v); }
__device__ __forceinline__ float hi32(u64 v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// ---- A: thread per hit point ------------------------------------------------------------------------------
__device__ __forceinline__ float fma_sat(float a, float b, float c) { float r; asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
// A variants that isolate what caps the loop: MODE 1 replaces the max by an FADD (FMA pipe only), MODE 2 by
// fma.sat + FADD (the any-survivor flag accumulated on the FMA pipe), MODE 3 = plain max but e_j = q_j - d_j.C
// with the warp-uniform sphere centre as the operand (uniform registers instead of per-thread tmp).
template <int NRV, int MODE>
__global__ void __launch_bounds__(128) k_Av(const float4 *__restrict__ sph, int nsph, const float4 *__restrict__ hp, int hp_per_thread,
                                            const float4 *__restrict__ rays, unsigned *out) {
  unsigned surv = 0;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  for (int it = 0; it < hp_per_thread; it++) {
    const float4 P = hp[(t + it * 977) & 0xffff];
    float dx[NRV], dy[NRV], dz[NRV], q[NRV];
    #pragma unroll
    for (int j = 0; j < NRV; j++) {
      const float4 r = rays[((t + it) * 17 + j) & 0xffff]; dx[j] = r.x; dy[j] = r.y; dz[j] = r.z;
      q[j] = __fmaf_rn(dz[j], P.z, __fmaf_rn(dy[j], P.y, dx[j] * P.x));
    }
    float4 a0 = sph[0], a1 = sph[1];
    #pragma unroll 1
    for (int s = 0; s < nsph; s += 2) {
      const float4 g0 = a0, g1 = a1;
      a0 = sph[s + 2]; a1 = sph[s + 3];
      const float x0 = P.x - g0.x, y0 = P.y - g0.y, z0 = P.z - g0.z;
      const float x1 = P.x - g1.x, y1 = P.y - g1.y, z1 = P.z - g1.z;
      const float T0 = __fmaf_rn(z0, z0, __fmaf_rn(y0, y0, x0 * x0));
      const float T1 = __fmaf_rn(z1, z1, __fmaf_rn(y1, y1, x1 * x1));
      const float k0 = __fmaf_rn(-EPS, T0 + g0.w, T0 - g0.w);
      const float k1 = __fmaf_rn(-EPS, T1 + g1.w, T1 - g1.w);
      float m0 = 0.f, m1 = 0.f;
      #pragma unroll
      for (int j = 0; j < NRV; j++) {
        if (MODE == 3) {
          const float e0 = __fmaf_rn(-dz[j], g0.z, __fmaf_rn(-dy[j], g0.y, __fmaf_rn(-dx[j], g0.x, q[j])));
          const float e1 = __fmaf_rn(-dz[j], g1.z, __fmaf_rn(-dy[j], g1.y, __fmaf_rn(-dx[j], g1.x, q[j])));
          m0 = fmaxf(m0, fabsf(e0)); m1 = fmaxf(m1, fabsf(e1));
        } else {
          const float e0 = __fmaf_rn(dz[j], z0, __fmaf_rn(dy[j], y0, dx[j] * x0));
          const float e1 = __fmaf_rn(dz[j], z1, __fmaf_rn(dy[j], y1, dx[j] * x1));
          if (MODE == 1) { m0 += e0; m1 += e1; }
          else if (MODE == 2) { m0 += fma_sat(e0, e0, -k0); m1 += fma_sat(e1, e1, -k1); }
          else { m0 = fmaxf(m0, fabsf(e0)); m1 = fmaxf(m1, fabsf(e1)); }
        }
      }
      if (MODE == 2) { surv += (m0 > 0.f) ? 1 : 0; surv += (m1 > 0.f) ? 1 : 0; }
      else { surv += (m0 * m0 >= k0) ? 1 : 0; surv += (m1 * m1 >= k1) ? 1 : 0; }
    }
  }
  if (surv) atomicAdd(out, surv);
}

// A4: sphere operands from __constant__ memory (warp-uniform index => uniform-register / constant-bank operands):
// e_j = q_j - d_j.C, q_j = d_j.P per ray; per (ray, sphere) three FFMAs whose second operand is not a register-file read.
__constant__ float4 c_sph[1100];
template <int NRV, int UNR>
__global__ void __launch_bounds__(128) k_A4(int nsph, const float4 *__restrict__ hp, int hp_per_thread,
                                            const float4 *__restrict__ rays, unsigned *out) {
  unsigned surv = 0;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  for (int it = 0; it < hp_per_thread; it++) {
    const float4 P = hp[(t + it * 977) & 0xffff];
    const float p2s = 1.9073486328125e-06f * __fmaf_rn(P.z, P.z, __fmaf_rn(P.y, P.y, P.x * P.x));
    float dx[NRV], dy[NRV], dz[NRV], q[NRV];
    #pragma unroll
    for (int j = 0; j < NRV; j++) {
      const float4 r = rays[((t + it) * 17 + j) & 0xffff]; dx[j] = -r.x; dy[j] = -r.y; dz[j] = -r.z;
      q[j] = __fmaf_rn(r.z, P.z, __fmaf_rn(r.y, P.y, r.x * P.x));
    }
    #pragma unroll 1
    for (int s = 0; s < nsph; s += UNR) {
      #pragma unroll
      for (int u = 0; u < UNR; u++) {
        const float4 g = c_sph[s + u];           // {cx, cy, cz, rho}: rho = r^2 (1 + eps) + 2^-19 |C|^2, host-side
        const float x = P.x - g.x, y = P.y - g.y, z = P.z - g.z;
        const float T = __fmaf_rn(z, z, __fmaf_rn(y, y, x * x));
        const float ka = __fmaf_rn(T, 1.0f - EPS, -g.w) - p2s;
        float m = 0.f;
        #pragma unroll
        for (int j = 0; j < NRV; j++)
          m = fmaxf(m, fabsf(__fmaf_rn(dz[j], g.z, __fmaf_rn(dy[j], g.y, __fmaf_rn(dx[j], g.x, q[j])))));
        surv += (m * m >= ka) ? 1 : 0;
      }
    }
  }
  if (surv) atomicAdd(out, surv);
}

// Raw issue-rate probes: N independent chains, REGS = how many distinct register-file operands an FFMA reads.
template <int MODE> __global__ void __launch_bounds__(128) k_rate(float *out, float ua, float ub, int iters) {
  float a[8], b[8], c[8];
  #pragma unroll
  for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 1e-3f + i; b[i] = 1.0f + i * 1e-4f + threadIdx.x * 1e-6f; c[i] = i; }
  for (int it = 0; it < iters; it++) {
    #pragma unroll
    for (int r = 0; r < 4; r++) {
      #pragma unroll
      for (int i = 0; i < 8; i++) {
        if (MODE == 0) c[i] = __fmaf_rn(a[i], b[i], c[i]);          // three distinct register operands
        if (MODE == 1) c[i] = __fmaf_rn(a[i], ua, c[i]);            // register, uniform, register
        if (MODE == 2) c[i] = __fmaf_rn(c[i], ua, ub);              // register, uniform, uniform
        if (MODE == 3) c[i] = __fmul_rn(c[i], b[i]);                // two registers
        if (MODE == 4) c[i] = __fmaf_rn(a[i], b[(i + 1) & 7], c[i]);  // three distinct registers, other pairing
        if (MODE == 5) c[i] = fmaxf(c[i], fmaxf(fabsf(a[i]), fabsf(b[i])));   // FMNMX3
        if (MODE == 6) c[i] = fmaxf(c[i], fabsf(a[i]));             // FMNMX
      }
    }
  }
  float sacc = 0;
  #pragma unroll
  for (int i = 0; i < 8; i++) sacc += c[i] + a[i] + b[i];
  if (sacc == 12345.f) out[0] = sacc;
}

template <int NRV>
__global__ void __launch_bounds__(128) k_A(const float4 *__restrict__ sph, int nsph, const float4 *__restrict__ hp, int hp_per_thread,
                                           const float4 *__restrict__ rays, unsigned *out) {
  unsigned surv = 0;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  for (int it = 0; it < hp_per_thread; it++) {
    const float4 P = hp[(t + it * 977) & 0xffff];
    float dx[NRV], dy[NRV], dz[NRV];
    #pragma unroll
    for (int j = 0; j < NRV; j++) { const float4 r = rays[((t + it) * 17 + j) & 0xffff]; dx[j] = r.x; dy[j] = r.y; dz[j] = r.z; }
    float4 a0 = sph[0], a1 = sph[1];
    #pragma unroll 1
    for (int s = 0; s < nsph; s += 2) {
      const float4 g0 = a0, g1 = a1;
      a0 = sph[s + 2]; a1 = sph[s + 3];
      const float x0 = P.x - g0.x, y0 = P.y - g0.y, z0 = P.z - g0.z;
      const float x1 = P.x - g1.x, y1 = P.y - g1.y, z1 = P.z - g1.z;
      const float T0 = __fmaf_rn(z0, z0, __fmaf_rn(y0, y0, x0 * x0));
      const float T1 = __fmaf_rn(z1, z1, __fmaf_rn(y1, y1, x1 * x1));
      const float k0 = __fmaf_rn(-EPS, T0 + g0.w, T0 - g0.w);
      const float k1 = __fmaf_rn(-EPS, T1 + g1.w, T1 - g1.w);
      float m0 = 0.f, m1 = 0.f;
      #pragma unroll
      for (int j = 0; j < NRV; j++) {
        m0 = fmaxf(m0, fabsf(__fmaf_rn(dz[j], z0, __fmaf_rn(dy[j], y0, dx[j] * x0))));
        m1 = fmaxf(m1, fabsf(__fmaf_rn(dz[j], z1, __fmaf_rn(dy[j], y1, dx[j] * x1))));
      }
      surv += (m0 * m0 >= k0) ? 1 : 0;
      surv += (m1 * m1 >= k1) ? 1 : 0;
    }
  }
  if (surv) atomicAdd(out, surv);
}

// ---- B / C: warp per hit point, lanes over spheres ---------------------------------------------------------
// shared memory per warp: rays of the current hit point as float4 {dx,dy,dz,-} (B) or duplicated pairs (C)
template <int S, bool PACKED>
__global__ void __launch_bounds__(128) k_B(const float4 *__restrict__ sph, int nsph, const float4 *__restrict__ hp, int hp_per_warp,
                                           const float4 *__restrict__ rays, int nrays_lo, int nrays_hi, unsigned *out) {
  __shared__ float4 rsm[4][2 * 18];   // per warp; packed layout uses two float4 per ray: {dx,dx,dy,dy} {dz,dz,-,-}
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int gw = blockIdx.x * 4 + w;
  unsigned surv = 0;
  for (int it = 0; it < hp_per_warp; it++) {
    const int h = (gw * 131 + it * 977) & 0xffff;
    const float4 P = hp[h];
    const int span = nrays_hi - nrays_lo + 1;
    int n = nrays_lo + (h % span);
    // stage the rays (lanes = rays), pad to an even count with a duplicate of the last ray
    __syncwarp();
    if (lane < 18) {
      const int j = lane < n ? lane : n - 1;
      const float4 r = rays[(h * 17 + j) & 0xffff];
      if (PACKED) { rsm[w][2 * lane] = make_float4(r.x, r.x, r.y, r.y); rsm[w][2 * lane + 1] = make_float4(r.z, r.z, 0.f, 0.f); }
      else rsm[w][lane] = r;
    }
    __syncwarp();
    n = (n + 1) & ~1;
    for (int c = 0; c < nsph; c += 32 * S) {
      if constexpr (!PACKED) {
        float x[S], y[S], z[S], ka[S], mx[S];
        #pragma unroll
        for (int s = 0; s < S; s++) {
          const float4 g = sph[c + s * 32 + lane];
          x[s] = P.x - g.x; y[s] = P.y - g.y; z[s] = P.z - g.z;
          const float T = __fmaf_rn(z[s], z[s], __fmaf_rn(y[s], y[s], x[s] * x[s]));
          ka[s] = __fmaf_rn(-EPS, T + g.w, T - g.w);
          mx[s] = 0.f;
        }
        #pragma unroll 1
        for (int j = 0; j < n; j += 2) {
          const float4 ra = rsm[w][j], rb = rsm[w][j + 1];
          #pragma unroll
          for (int s = 0; s < S; s++) {
            const float ea = __fmaf_rn(ra.z, z[s], __fmaf_rn(ra.y, y[s], ra.x * x[s]));
            const float eb = __fmaf_rn(rb.z, z[s], __fmaf_rn(rb.y, y[s], rb.x * x[s]));
            mx[s] = fmaxf(mx[s], fmaxf(fabsf(ea), fabsf(eb)));
          }
        }
        float v = -1.f;
        #pragma unroll
        for (int s = 0; s < S; s++) v = fmaxf(v, __fmaf_rn(mx[s], mx[s], -ka[s]));
        if (v >= 0.f) {
          #pragma unroll
          for (int s = 0; s < S; s++) surv += (mx[s] * mx[s] >= ka[s]) ? 1 : 0;
        }
      } else {
        static_assert(!PACKED || (S % 2 == 0), "pairs");
        u64 x[S / 2], y[S / 2], z[S / 2], ka[S / 2];
        float mx[S];
        const u64 Px = pk(P.x, P.x), Py = pk(P.y, P.y), Pz = pk(P.z, P.z), ME = pk(-EPS, -EPS);
        #pragma unroll
        for (int s = 0; s < S / 2; s++) {
          const float4 g0 = sph[c + (2 * s) * 32 + lane], g1 = sph[c + (2 * s + 1) * 32 + lane];
          x[s] = sub2(Px, pk(g0.x, g1.x)); y[s] = sub2(Py, pk(g0.y, g1.y)); z[s] = sub2(Pz, pk(g0.z, g1.z));
          const u64 T = fma2(z[s], z[s], fma2(y[s], y[s], mul2(x[s], x[s])));
          const u64 R = pk(g0.w, g1.w);
          ka[s] = fma2(ME, add2(T, R), sub2(T, R));
          mx[2 * s] = 0.f; mx[2 * s + 1] = 0.f;
        }
        #pragma unroll 1
        for (int j = 0; j < n; j += 2) {
          const float4 a0 = rsm[w][2 * j], a1 = rsm[w][2 * j + 1], b0 = rsm[w][2 * j + 2], b1 = rsm[w][2 * j + 3];
          const u64 ax = pk(a0.x, a0.y), ay = pk(a0.z, a0.w), az = pk(a1.x, a1.y);
          const u64 bx = pk(b0.x, b0.y), by = pk(b0.z, b0.w), bz = pk(b1.x, b1.y);
          #pragma unroll
          for (int s = 0; s < S / 2; s++) {
            const u64 ea = fma2(az, z[s], fma2(ay, y[s], mul2(ax, x[s])));
            const u64 eb = fma2(bz, z[s], fma2(by, y[s], mul2(bx, x[s])));
            mx[2 * s] = fmaxf(mx[2 * s], fmaxf(fabsf(lo32(ea)), fabsf(lo32(eb))));
            mx[2 * s + 1] = fmaxf(mx[2 * s + 1], fmaxf(fabsf(hi32(ea)), fabsf(hi32(eb))));
          }
        }
        float v = -1.f;
        #pragma unroll
        for (int s = 0; s < S / 2; s++) {
          const u64 m2 = pk(mx[2 * s], mx[2 * s + 1]);
          const u64 q = fma2(m2, m2, ka[s] ^ 0x8000000080000000ull);
          v = fmaxf(v, fmaxf(lo32(q), hi32(q)));
        }
        if (v >= 0.f) {
          #pragma unroll
          for (int s = 0; s < S / 2; s++) {
            surv += (mx[2 * s] * mx[2 * s] >= lo32(ka[s])) ? 1 : 0;
            surv += (mx[2 * s + 1] * mx[2 * s + 1] >= hi32(ka[s])) ? 1 : 0;
          }
        }
      }
    }
  }
  if (surv) atomicAdd(out, surv);
}

static float frand() { return (float)rand() / (float)RAND_MAX; }

template <typename F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int k = 0; k < 3; k++) {
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  return best;
}

int main(int argc, char **argv) {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sm = p.multiProcessorCount;
  const double clk = 1.965e9;
  const int nsph = 1024;
  std::vector<float4> hs(nsph + 64), hh(65536), hr(65536);
  srand(7);
  for (int i = 0; i < nsph + 64; i++) { float r = 0.2f + 0.5f * frand(); hs[i] = make_float4(40 * frand() - 20, 10 * frand(), -10 - 40 * frand(), i < nsph ? r * r : -INFINITY); }
  for (auto &h : hh) h = make_float4(40 * frand() - 20, 10 * frand(), -10 - 40 * frand(), 0.f);
  for (auto &r : hr) { float x = frand() - .5f, y = frand() - .5f, z = frand() - .5f, l = sqrtf(x * x + y * y + z * z) + 1e-9f; r = make_float4(x / l, y / l, z / l, 0.f); }
  float4 *ds, *dh, *dr; unsigned *dout;
  cudaMalloc(&ds, hs.size() * 16); cudaMalloc(&dh, hh.size() * 16); cudaMalloc(&dr, hr.size() * 16); cudaMalloc(&dout, 4);
  cudaMemcpy(ds, hs.data(), hs.size() * 16, cudaMemcpyHostToDevice);
  cudaMemcpy(dh, hh.data(), hh.size() * 16, cudaMemcpyHostToDevice);
  cudaMemcpy(dr, hr.data(), hr.size() * 16, cudaMemcpyHostToDevice);
  cudaMemset(dout, 0, 4);
  auto report = [&](const char *name, int cfg, int nr_lo, int nr_hi, double hps, float ms, int ctas_per_sm) {
    const double nr = 0.5 * (nr_lo + nr_hi);
    const double tests = hps * nsph * nr;
    const double fma_instr = hps * nsph * (9.0 + 3.0 * nr) / 32.0;   // warp instructions, scalar-equivalent
    const double pipe = fma_instr / (ms * 1e-3) / (sm * 4.0 * clk);
    const double alg = hps * nsph * (10.0 + 8.0 * nr) / (ms * 1e-3) / 1e12;
    printf("{\"shape\": \"%s\", \"cfg\": %d, \"rays\": [%d, %d], \"ctas_per_sm\": %d, \"ms\": %.3f, \"Gtests_s\": %.1f, \"fma_pipe_equiv\": %.3f, \"alg_tflops\": %.2f, \"alg_frac\": %.3f},\n",
           name, cfg, nr_lo, nr_hi, ctas_per_sm, ms, tests / ms / 1e6, pipe, alg, alg / (sm * 128 * 2 * clk / 1e12));
    fflush(stdout);
  };
  if (argc > 1) {   // one configuration, for ncu:  scan_shapes <A0|A1|A2|A3|B|C> <cps>
    const char *w = argv[1];
    const int cps = argc > 2 ? atoi(argv[2]) : 4;
    const int g = sm * cps;
    float ms = 0;
    printf("[\n");
    if (w[0] == 'A') {
      const int hpt = 24; const double hps = (double)g * 128 * hpt;
      if (w[1] == '0') ms = timeit([&] { k_Av<17, 0><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); });
      if (w[1] == '1') ms = timeit([&] { k_Av<17, 1><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); });
      if (w[1] == '2') ms = timeit([&] { k_Av<17, 2><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); });
      if (w[1] == '3') ms = timeit([&] { k_Av<17, 3><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); });
      if (w[1] == '4') {
        std::vector<float4> hc(1100, make_float4(0.f, 0.f, 0.f, INFINITY));
        for (int i = 0; i < nsph; i++) { const float4 q = hs[i]; hc[i] = make_float4(q.x, q.y, q.z, q.w * (1.0f + EPS) + 1.9073486328125e-06f * (q.x * q.x + q.y * q.y + q.z * q.z)); }
        cudaMemcpyToSymbol(c_sph, hc.data(), 1100 * 16);
        ms = timeit([&] { k_A4<17, 2><<<g, 128>>>(nsph, dh, hpt, dr, dout); });
      }
      report(w, 17, 17, 17, hps, ms, cps);
    } else {
      const int hpw = 400; const double hps = (double)g * 4 * hpw;
      if (w[0] == 'B') ms = timeit([&] { k_B<8, false><<<g, 128>>>(ds, nsph, dh, hpw, dr, 16, 16, dout); });
      else ms = timeit([&] { k_B<8, true><<<g, 128>>>(ds, nsph, dh, hpw, dr, 16, 16, dout); });
      report(w, 8, 16, 16, hps, ms, cps);
    }
    printf("{}]\n");
    return 0;
  }
  printf("[\n");
  for (int cps : {4}) {
    const int g = sm * cps, hpt = 24;
    const double hps = (double)g * 128 * hpt;
    float ms;
    ms = timeit([&] { k_Av<17, 0><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av0 max3", 17, 17, 17, hps, ms, cps);
    ms = timeit([&] { k_Av<17, 1><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av1 fadd instead of max", 17, 17, 17, hps, ms, cps);
    ms = timeit([&] { k_Av<17, 2><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av2 fma.sat + fadd", 17, 17, 17, hps, ms, cps);
    ms = timeit([&] { k_Av<17, 3><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av3 uniform centre operand", 17, 17, 17, hps, ms, cps);
    ms = timeit([&] { k_Av<9, 0><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av0 max3", 9, 9, 9, hps, ms, cps);
    ms = timeit([&] { k_Av<9, 1><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av1 fadd instead of max", 9, 9, 9, hps, ms, cps);
    ms = timeit([&] { k_Av<9, 2><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av2 fma.sat + fadd", 9, 9, 9, hps, ms, cps);
    ms = timeit([&] { k_Av<9, 3><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("Av3 uniform centre operand", 9, 9, 9, hps, ms, cps);
  }
  {
    std::vector<float4> hc(1100, make_float4(0.f, 0.f, 0.f, INFINITY));
    for (int i = 0; i < nsph; i++) {
      const float4 g = hs[i];
      hc[i] = make_float4(g.x, g.y, g.z, g.w * (1.0f + EPS) + 1.9073486328125e-06f * (g.x * g.x + g.y * g.y + g.z * g.z));
    }
    cudaMemcpyToSymbol(c_sph, hc.data(), 1100 * 16);
    float *fo; cudaMalloc(&fo, 4);
    for (int cps : {3, 4}) {
      const int g = sm * cps, hpt = 24;
      const double hps = (double)g * 128 * hpt;
      float ms;
      ms = timeit([&] { k_A4<17, 1><<<g, 128>>>(nsph, dh, hpt, dr, dout); }); report("A4 const-bank centre, 1 sphere/trip", 17, 17, 17, hps, ms, cps);
      ms = timeit([&] { k_A4<17, 2><<<g, 128>>>(nsph, dh, hpt, dr, dout); }); report("A4 const-bank centre, 2 spheres/trip", 17, 17, 17, hps, ms, cps);
      ms = timeit([&] { k_A4<17, 4><<<g, 128>>>(nsph, dh, hpt, dr, dout); }); report("A4 const-bank centre, 4 spheres/trip", 17, 17, 17, hps, ms, cps);
      ms = timeit([&] { k_A4<9, 2><<<g, 128>>>(nsph, dh, hpt, dr, dout); }); report("A4 const-bank centre, 2 spheres/trip", 9, 9, 9, hps, ms, cps);
      ms = timeit([&] { k_A4<9, 4><<<g, 128>>>(nsph, dh, hpt, dr, dout); }); report("A4 const-bank centre, 4 spheres/trip", 9, 9, 9, hps, ms, cps);
      ms = timeit([&] { k_A4<5, 4><<<g, 128>>>(nsph, dh, hpt, dr, dout); }); report("A4 const-bank centre, 4 spheres/trip", 5, 5, 5, hps, ms, cps);
    }
    const char *names[7] = {"FFMA 3 distinct regs", "FFMA reg,uniform,reg", "FFMA reg,uniform,uniform", "FMUL 2 regs", "FFMA 3 regs (other pairing)", "FMNMX3 |a| |b|", "FMNMX |a|"};
    for (int wps : {1, 2, 4, 8}) {
      const int g = sm * wps, iters = 4000;
      float ms[7];
      ms[0] = timeit([&] { k_rate<0><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      ms[1] = timeit([&] { k_rate<1><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      ms[2] = timeit([&] { k_rate<2><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      ms[3] = timeit([&] { k_rate<3><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      ms[4] = timeit([&] { k_rate<4><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      ms[5] = timeit([&] { k_rate<5><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      ms[6] = timeit([&] { k_rate<6><<<g, 128>>>(fo, 1.0001f, 0.5f, iters); });
      for (int k = 0; k < 7; k++)
        printf("{\"probe\": \"%s\", \"warps_per_smsp\": %d, \"inst_per_clk_smsp\": %.3f},\n", names[k], wps,
               (double)g * 4 * iters * 32.0 / (ms[k] * 1e-3) / (sm * 4.0 * clk));
    }
  }
  if (getenv("SCAN_SHAPES_SHORT")) { printf("{}]\n"); return 0; }
  // A: 128 threads per CTA, 4 CTAs/SM resident (round-1 occupancy) and 3
  for (int cps : {3, 4}) {
    const int g = sm * cps, hpt = 24;
    const double hps = (double)g * 128 * hpt;
    float ms;
    ms = timeit([&] { k_A<17><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("A thread/hp NRV", 17, 17, 17, hps, ms, cps);
    ms = timeit([&] { k_A<9><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("A thread/hp NRV", 9, 9, 9, hps, ms, cps);
    ms = timeit([&] { k_A<5><<<g, 128>>>(ds, nsph, dh, hpt, dr, dout); }); report("A thread/hp NRV", 5, 5, 5, hps, ms, cps);
  }
  for (int cps : {2, 4, 6, 8}) {
    const int g = sm * cps, hpw = 400;
    const double hps = (double)g * 4 * hpw;
    float ms;
    const int lohi[4][2] = {{16, 16}, {8, 8}, {4, 4}, {2, 14}};
    for (auto &lh : lohi) {
      ms = timeit([&] { k_B<4, false><<<g, 128>>>(ds, nsph, dh, hpw, dr, lh[0], lh[1], dout); }); report("B lanes=spheres scalar S", 4, lh[0], lh[1], hps, ms, cps);
      ms = timeit([&] { k_B<8, false><<<g, 128>>>(ds, nsph, dh, hpw, dr, lh[0], lh[1], dout); }); report("B lanes=spheres scalar S", 8, lh[0], lh[1], hps, ms, cps);
      ms = timeit([&] { k_B<4, true><<<g, 128>>>(ds, nsph, dh, hpw, dr, lh[0], lh[1], dout); }); report("C lanes=spheres f32x2 S", 4, lh[0], lh[1], hps, ms, cps);
      ms = timeit([&] { k_B<8, true><<<g, 128>>>(ds, nsph, dh, hpw, dr, lh[0], lh[1], dout); }); report("C lanes=spheres f32x2 S", 8, lh[0], lh[1], hps, ms, cps);
      ms = timeit([&] { k_B<16, true><<<g, 128>>>(ds, nsph, dh, hpw, dr, lh[0], lh[1], dout); }); report("C lanes=spheres f32x2 S", 16, lh[0], lh[1], hps, ms, cps);
    }
  }
  unsigned o = 0; cudaMemcpy(&o, dout, 4, cudaMemcpyDeviceToHost);
  printf("{\"survivors\": %u, \"err\": \"%s\"}]\n", o, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
