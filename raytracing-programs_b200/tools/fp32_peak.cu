// tools/fp32_peak.cu -- FP32 pipe calibration for the roofline (SURVEY.md section 8d asks for an
// FFMA microbenchmark on the box).  Measures scalar FFMA, packed FFMA2 (fma.rn.f32x2, sm_100+)
// and the instruction shape of the sphere reject filter (FMUL + 3 FFMA + 1/2 FMNMX3 per test).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 pack(float lo, float hi) { return ((u64)__float_as_uint(hi) << 32) | __float_as_uint(lo); }

template <int N> __global__ void k_ffma(float *out, float a, float b, int iters) {
  float x[N];
  #pragma unroll
  for (int i = 0; i < N; i++) x[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < iters; it++) {
    #pragma unroll
    for (int i = 0; i < N; i++) x[i] = __fmaf_rn(x[i], a, b);
  }
  float s = 0; 
  #pragma unroll
  for (int i = 0; i < N; i++) s += x[i];
  if (s == 12345.f) out[0] = s;
}
template <int N> __global__ void k_ffma2(float *out, float a, float b, int iters) {
  u64 x[N]; u64 A = pack(a, a), B = pack(b, b);
  #pragma unroll
  for (int i = 0; i < N; i++) x[i] = pack(threadIdx.x * 1e-3f + i, i);
  for (int it = 0; it < iters; it++) {
    #pragma unroll
    for (int i = 0; i < N; i++) x[i] = ffma2(x[i], A, B);
  }
  u64 s = 0;
  #pragma unroll
  for (int i = 0; i < N; i++) s ^= x[i];
  if (s == 12345) out[0] = 1.f;
}
// filter shape: per (ray, sphere): e = fma(dz,z,fma(dy,y,dx*x)); q = fma(e,e,k); m = max(m,q)
template <int NR> __global__ void k_filter(float *out, const float4 *sph, int nsph, int iters) {
  float dx[NR], dy[NR], dz[NR];
  #pragma unroll
  for (int j = 0; j < NR; j++) { dx[j] = threadIdx.x * 1e-3f + j; dy[j] = j * 0.5f; dz[j] = 1.f - j; }
  float m = -1.f;
  for (int it = 0; it < iters; it++)
    for (int s = 0; s < nsph; s++) {
      float4 h = sph[s];
      #pragma unroll
      for (int j = 0; j < NR; j++) {
        float e = __fmaf_rn(dz[j], h.z, __fmaf_rn(dy[j], h.y, dx[j] * h.x));
        m = fmaxf(m, __fmaf_rn(e, e, h.w));
      }
    }
  if (m == 12345.f) out[0] = m;
}
template <int NR> __global__ void k_filter2(float *out, const float4 *sph, int nsph, int iters) {
  // packed: two spheres per FFMA2 (lo = sphere s, hi = sphere s+1), ray components duplicated
  u64 dx[NR], dy[NR], dz[NR];
  #pragma unroll
  for (int j = 0; j < NR; j++) { float a = threadIdx.x * 1e-3f + j, b = j * 0.5f, c = 1.f - j; dx[j] = pack(a, a); dy[j] = pack(b, b); dz[j] = pack(c, c); }
  float m = -1.f;
  for (int it = 0; it < iters; it++)
    for (int s = 0; s < nsph; s += 2) {
      float4 h0 = sph[s], h1 = sph[s + 1];
      u64 X = pack(h0.x, h1.x), Y = pack(h0.y, h1.y), Z = pack(h0.z, h1.z), K = pack(h0.w, h1.w);
      #pragma unroll
      for (int j = 0; j < NR; j++) {
        u64 e = ffma2(dz[j], Z, ffma2(dy[j], Y, ffma2(dx[j], X, 0ull)));
        u64 q = ffma2(e, e, K);
        m = fmaxf(m, fmaxf(__uint_as_float((unsigned)q), __uint_as_float((unsigned)(q >> 32))));
      }
    }
  if (m == 12345.f) out[0] = m;
}
// packed over RAY PAIRS: lo = ray 2j, hi = ray 2j+1; sphere operands duplicated into both halves
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
template <int NP> __global__ void k_filter3(float *out, const float4 *sph, int nsph, int iters, float px, float py, float pz) {
  u64 dx[NP], dy[NP], dz[NP];
  #pragma unroll
  for (int j = 0; j < NP; j++) { float a = threadIdx.x * 1e-3f + j, b = j * 0.5f, c = 1.f - j; dx[j] = pack(a, a + 1); dy[j] = pack(b, b + 1); dz[j] = pack(c, c + 1); }
  float m = -1.f;
  for (int it = 0; it < iters; it++)
    for (int s = 0; s < nsph; s++) {
      float4 h = sph[s];
      float x = px - h.x, y = py - h.y, z = pz - h.z;
      float T = __fmaf_rn(z, z, __fmaf_rn(y, y, x * x));
      float k = __fmaf_rn(3.8e-6f, T + h.w, h.w - T);
      u64 X = pack(x, x), Y = pack(y, y), Z = pack(z, z), K = pack(k, k);
      #pragma unroll
      for (int j = 0; j < NP; j++) {
        u64 e = ffma2(dz[j], Z, ffma2(dy[j], Y, fmul2(dx[j], X)));
        u64 q = ffma2(e, e, K);
        m = fmaxf(m, fmaxf(__uint_as_float((unsigned)q), __uint_as_float((unsigned)(q >> 32))));
      }
    }
  if (m == 12345.f) out[0] = m;
}
template <typename F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main() {
  float *out; cudaMalloc(&out, 4);
  float4 *sph; cudaMalloc(&sph, 1024 * 16); cudaMemset(sph, 0, 1024 * 16);
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sm = p.multiProcessorCount;
  printf("{\"sms\": %d, \"results\": [\n", sm);
  for (int tpb : {128, 256, 512, 1024}) {
    int blocks = sm * (2048 / tpb) / 2;   // half of max occupancy: 1024 threads/SM = 8 warps/SMSP ... 
    for (int occ = 1; occ <= 2; occ++) {
      int g = sm * occ * (tpb >= 1024 ? 1 : 1024 / tpb);
      if (occ == 2 && tpb == 1024) g = sm * 2;
      const int it = 20000;
      float ms = timeit([&] { k_ffma<16><<<g, tpb>>>(out, 1.0001f, 0.5f, it); });
      double tf = 2.0 * 16 * it * (double)g * tpb / ms / 1e9;
      float ms2 = timeit([&] { k_ffma2<8><<<g, tpb>>>(out, 1.0001f, 0.5f, it); });
      double tf2 = 4.0 * 8 * it * (double)g * tpb / ms2 / 1e9;
      printf("{\"threads_per_sm\": %d, \"tpb\": %d, \"ffma_tflops\": %.2f, \"ffma2_tflops\": %.2f},\n", g / sm * tpb, tpb, tf, tf2);
    }
    (void)blocks;
  }
  for (int wps : {1, 2, 4, 8}) {   // warps per SMSP
    int tpb = 128, g = sm * wps;
    float ms = timeit([&] { k_filter<17><<<g, tpb>>>(out, sph, 1024, 40); });
    float ms2 = timeit([&] { k_filter2<17><<<g, tpb>>>(out, sph, 1024, 40); });
    float ms3 = timeit([&] { k_filter3<9><<<g, tpb>>>(out, sph, 1024, 40, 0.1f, 0.2f, 0.3f); });
    double tests = 17.0 * 1024 * 40 * (double)g * tpb;
    printf("{\"warps_per_smsp\": %d, \"filter3_raypair_Gtests_s\": %.1f},\n", wps, tests / 17 * 18 / ms3 / 1e6);
    printf("{\"warps_per_smsp\": %d, \"filter_Gtests_s\": %.1f, \"filter_instr_per_clk_smsp\": %.3f, \"filter2_Gtests_s\": %.1f},\n",
           wps, tests / ms / 1e6, tests * 4.5 / 32 / (ms * 1e-3) / (sm * 4) / (p.clockRate * 1e3), tests / ms2 / 1e6);
  }
  printf("{}]}\n");
  return 0;
}
