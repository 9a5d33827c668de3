// csrc/rt_p3.cuh -- the body of ppm_WriteOutP3 (C/ppm.c:168-184) on the GPU: every byte of the
// RGB8 frame becomes its decimal text followed by '\n' (2..4 bytes), the texts concatenated in
// pixmap order.  HBM-bound; output positions are a prefix sum over variable lengths:
//
//   k_p3_count  text length of every span of 65536 values (streaming read of the frame,
//               byte-parallel compares, one atomicAdd per 4096 values)
//   k_p3_scan   exclusive scan of the span lengths -> 64-bit start offsets (one CTA; a
//               32768^2 frame has 49 152 spans)
//   k_p3_emit   one CTA per span, 16 chunks of 4096 values each: format, stage in shared
//               memory, store 16 bytes at a time.  No CTA waits for another one.
//
//   algorithmic traffic per value: 1 byte read + len(value) bytes written (2..4); the count
//   pass reads the frame a second time (1 more byte per value of DRAM traffic).
//   (A single-pass variant with a decoupled look-back over 4096-value chunks was measured
//   first: 2.7 ms for the 16384^2 frame, bound by the look-back latency chain, not by HBM.)
//
// k_p3_emit, per thread: one 16-byte load = 16 values (the next chunk's load is in flight while
// this one is formatted); text of a value from a 256-entry table replicated once per lane in
// shared memory (entry v of lane l at word 32*v + l: bank = lane, so the table reads of a warp
// never conflict whatever the pixel values are); the 16 texts are concatenated in a register
// shift accumulator and leave as aligned 32-bit shared-memory stores (only the first and the
// last word of a thread can be shared with a neighbour: those are OR-ed atomically into the
// zeroed buffer).  The chunk is staged at the 16-byte phase of its place in the output, so it
// leaves as aligned 16-byte streaming stores; bytes only at the two ragged ends.
#pragma once
#include <stdint.h>

#define P3_THREADS 256
#define P3_VPT 16
#define P3_CHUNK (P3_THREADS * P3_VPT)            /* values per chunk (one block iteration) */
#define P3_SPAN_CHUNKS 16
#define P3_SPAN (P3_CHUNK * P3_SPAN_CHUNKS)       /* values per span (one entry of the offset table) */
#define P3_TXT_WORDS 4104                         /* 15 + 4*4096 bytes, rounded up to 16 */
#define P3_LUT_WORDS (256 * 32)
#define P3_SMEM_BYTES ((P3_LUT_WORDS + P3_TXT_WORDS) * 4)

// "<decimal>\n" of v as little-endian bytes of one word ("7\n" = 0x00000a37, "255\n" = 0x0a353532)
__host__ __device__ inline uint32_t p3_word(unsigned v) {
  const unsigned h = v / 100u, t = (v / 10u) % 10u, o = v % 10u;
  if (v >= 100u) return (0x30u + h) | ((0x30u + t) << 8) | ((0x30u + o) << 16) | (0x0au << 24);
  if (v >= 10u) return (0x30u + t) | ((0x30u + o) << 8) | (0x0au << 16);
  return (0x30u + o) | (0x0au << 8);
}

// text bytes of the four values packed in one word: 2 each, +1 per value > 9, +1 per value > 99
__device__ __forceinline__ unsigned p3_len4(uint32_t x) {
  return 8u + (__popc(__vcmpgtu4(x, 0x09090909u)) >> 3) + (__popc(__vcmpgtu4(x, 0x63636363u)) >> 3);
}

// 16 values of thread `tid` of chunk `chunk` (absent values read as none: nv < 16 only in the last chunk)
__device__ __forceinline__ void p3_load16(const uint8_t *__restrict__ rgb, unsigned long long n_values,
                                          unsigned long long chunk, int tid, uint4 &q, int &nv) {
  q = make_uint4(0u, 0u, 0u, 0u);
  nv = 0;
  const unsigned long long v0 = chunk * P3_CHUNK + (unsigned)tid * P3_VPT;
  if (v0 + P3_VPT <= n_values) {
    q = __ldcs(reinterpret_cast<const uint4 *>(rgb + v0));
    nv = P3_VPT;
  } else if (v0 < n_values) {
    nv = (int)(n_values - v0);
    uint32_t r[4] = {0u, 0u, 0u, 0u};
    for (int k = 0; k < nv; k++) r[k >> 2] |= (uint32_t)rgb[v0 + k] << (8 * (k & 3));
    q = make_uint4(r[0], r[1], r[2], r[3]);
  }
}

// ---------------------------------------------------------------------------------
// k_p3_count: span_len[s] += text bytes of the values of span s (span_len zeroed before).
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(P3_THREADS)
k_p3_count(const uint8_t *__restrict__ rgb, unsigned long long n_values, unsigned long long n_chunks,
           unsigned *__restrict__ span_len) {
  __shared__ unsigned warp_tot[P3_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (unsigned long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    uint4 q;
    int nv;
    p3_load16(rgb, n_values, chunk, tid, q, nv);
    unsigned len;
    if (nv == P3_VPT) {
      len = p3_len4(q.x) + p3_len4(q.y) + p3_len4(q.z) + p3_len4(q.w);
    } else {
      const uint32_t r[4] = {q.x, q.y, q.z, q.w};
      len = 0;
      for (int k = 0; k < nv; k++) { const unsigned v = (r[k >> 2] >> (8 * (k & 3))) & 0xffu; len += 2u + (v > 9u) + (v > 99u); }
    }
    len = __reduce_add_sync(0xffffffffu, len);
    if (lane == 0) warp_tot[warp] = len;
    __syncthreads();
    if (tid == 0) {
      unsigned L = 0;
      #pragma unroll
      for (int k = 0; k < P3_THREADS / 32; k++) L += warp_tot[k];
      atomicAdd(&span_len[chunk / P3_SPAN_CHUNKS], L);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------
// k_p3_scan (one CTA of 1024 threads): span_off[s] = base_off + sum of span_len[< s];
// span_off[n_spans] = end offset of the whole text.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
k_p3_scan(const unsigned *__restrict__ span_len, unsigned long long *__restrict__ span_off, unsigned n_spans,
          unsigned long long base_off) {
  __shared__ unsigned long long warp_tot[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned per = (n_spans + 1023u) / 1024u;
  const unsigned lo = min(n_spans, tid * per), hi = min(n_spans, lo + per);
  unsigned long long mine = 0;
  for (unsigned i = lo; i < hi; i++) mine += span_len[i];
  unsigned long long inc = mine;
  #pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const unsigned long long u = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += u;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    unsigned long long t = warp_tot[lane], ti = t;
    #pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned long long u = __shfl_up_sync(0xffffffffu, ti, off);
      if (lane >= off) ti += u;
    }
    warp_tot[lane] = ti - t;   // exclusive prefix of the warp totals
  }
  __syncthreads();
  unsigned long long run = base_off + warp_tot[warp] + inc - mine;
  for (unsigned i = lo; i < hi; i++) { span_off[i] = run; run += span_len[i]; }
  if (tid == 1023) span_off[n_spans] = run;
}

// ---------------------------------------------------------------------------------
// k_p3_emit: CTA s formats span s at text + span_off[s].  text 16-byte aligned.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(P3_THREADS, 4)
k_p3_emit(const uint8_t *__restrict__ rgb, unsigned long long n_values, unsigned long long n_chunks,
          const unsigned long long *__restrict__ span_off, unsigned n_spans, char *__restrict__ text) {
  extern __shared__ __align__(16) uint32_t p3_smem[];
  uint32_t *lut = p3_smem;
  uint32_t *txt = p3_smem + P3_LUT_WORDS;
  __shared__ unsigned warp_tot[2][P3_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < P3_LUT_WORDS; i += P3_THREADS) lut[i] = p3_word((unsigned)i >> 5);
  for (int i = tid; i < P3_TXT_WORDS; i += P3_THREADS) txt[i] = 0u;
  const uint32_t *my_lut = lut + lane;
  __syncthreads();
  for (unsigned span = blockIdx.x; span < n_spans; span += gridDim.x) {
    unsigned long long B = span_off[span];                     // running output offset inside the span
    const unsigned long long c0 = (unsigned long long)span * P3_SPAN_CHUNKS;
    const int nc = (int)min((unsigned long long)P3_SPAN_CHUNKS, n_chunks - c0);
    uint4 cur;
    int cur_nv;
    p3_load16(rgb, n_values, c0, tid, cur, cur_nv);
    for (int c = 0; c < nc; c++) {
      uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
      int nxt_nv = 0;
      if (c + 1 < nc) p3_load16(rgb, n_values, c0 + c + 1, tid, nxt, nxt_nv);   // in flight during this chunk
      // ---- 16 values -> 16 text words ---------------------------------------------
      const bool ragged = (c0 + c + 1) * P3_CHUNK > n_values;    // block-uniform: only the frame's last chunk
      const uint32_t raw[4] = {cur.x, cur.y, cur.z, cur.w};
      uint32_t w[P3_VPT];
      unsigned bits = 0;
      #pragma unroll
      for (int k = 0; k < P3_VPT; k++) {
        const unsigned v = (raw[k >> 2] >> (8 * (k & 3))) & 0xffu;
        w[k] = my_lut[v << 5];
        if (ragged && k >= cur_nv) w[k] = 0u;
        bits += 32u - ((unsigned)__clz(w[k]) & 0x38u);          // 8 * text length; 0 for an absent value
      }
      // ---- exclusive scan of the thread lengths inside the chunk -----------------------
      const unsigned tb = bits >> 3;
      unsigned inc = tb;
      #pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const unsigned u = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += u;
      }
      if (lane == 31) warp_tot[c & 1][warp] = inc;
      __syncthreads();                                         // (1) warp totals visible; staging buffer is zero
      unsigned before = 0, L = 0;
      #pragma unroll
      for (int k = 0; k < P3_THREADS / 32; k++) { const unsigned t = warp_tot[c & 1][k]; before += k < warp ? t : 0u; L += t; }
      const unsigned a = (unsigned)(B & 15ull);
      // ---- concatenate the 16 texts into the staged chunk -----------------------------
      {
        const unsigned o = a + before + inc - tb;
        uint32_t *wp = txt + (o >> 2);
        unsigned fill = (o & 3u) * 8u;
        uint32_t lo = 0u;
        #pragma unroll
        for (int k = 0; k < P3_VPT; k++) {
          const uint32_t t_hi = __funnelshift_l(w[k], 0u, fill); // bits that spill into the next word
          lo |= w[k] << fill;
          fill += 32u - ((unsigned)__clz(w[k]) & 0x38u);
          if (fill >= 32u) {
            // the first word of a thread may hold a neighbour's bytes: it is complete after at
            // most two values (>= 2 bytes each), so only k < 2 needs the atomic
            if (k < 2) atomicOr(wp, lo); else *wp = lo;
            wp++;
            lo = t_hi;
          }
          fill &= 31u;
        }
        if (fill) atomicOr(wp, lo);
      }
      __syncthreads();                                         // (2) chunk staged
      // ---- staged text -> global, 16 bytes at a time, and re-zero the staging buffer ------
      // Groups that are complete leave as one 16-byte store.  The ragged last group of a chunk
      // is not stored: thread 0 moves it to the front of the staging buffer, where the next
      // chunk's text continues it (same 16-byte phase), so only the first group of a span
      // (its head belongs to the previous span) and the last group of the span's last chunk
      // take the byte path.
      {
        char *g0 = text + (B - a);
        const unsigned end = a + L, ngroups = (end + 15u) >> 4;
        const unsigned head = c > 0 ? 0u : a;                   // first byte of the staging buffer that is this span's
        const bool last_chunk = c + 1 == nc;
        uint4 *t4 = reinterpret_cast<uint4 *>(txt);
        auto put = [&](unsigned g, uint4 q) {
          const unsigned lo_b = g << 4, hi_b = lo_b + 16u;
          if (lo_b >= head && hi_b <= end) {
            __stcs(reinterpret_cast<uint4 *>(g0) + g, q);
          } else if (hi_b <= end || last_chunk) {
            const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
            #pragma unroll 1
            for (unsigned b = 0; b < 16u; b++)
              if (lo_b + b >= head && lo_b + b < end) g0[lo_b + b] = (char)((qq[b >> 2] >> (8u * (b & 3u))) & 0xffu);
          }
        };
        for (unsigned g = tid; g + 1 < ngroups; g += P3_THREADS) {
          const uint4 q = t4[g];
          t4[g] = make_uint4(0u, 0u, 0u, 0u);
          put(g, q);
        }
        if (tid == 0 && ngroups > 0) {                          // after its own groups (group 0 among them)
          const uint4 q = t4[ngroups - 1];
          t4[ngroups - 1] = make_uint4(0u, 0u, 0u, 0u);
          put(ngroups - 1, q);
          if ((end & 15u) && !last_chunk) t4[0] = q;            // carry: bytes [0, end & 15) of the next chunk's first group
        }
      }
      B += L;
      cur = nxt;
      cur_nv = nxt_nv;
      // no third barrier: the next chunk writes the staging buffer only after its barrier (1),
      // which every thread reaches after finishing this copy-out; warp_tot is double-buffered
    }
    __syncthreads();
  }
}
