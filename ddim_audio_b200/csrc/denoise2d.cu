// denoise_2d post-filter (reference External/UPU/signal/denoise.py:45-70, called on every returned x0 at
// runners/diffusion.py:400-401) on x[B][2][T][F] fp32 -- SURVEY.md section 8 row f1.
//
// The reference materialises r five times over (two medians through full sorts, then per iteration a norm, a division, two
// means and two broadcast subtractions: ~30 launches and ~40 tensor passes).  Every update it makes to r is a per-row
// (b, c, t) or per-column (b, c, f) offset, so here r is never written until the end: the offsets are kept per step,
//     r_k = ((((x - rowoff_0[t]) - coloff_0[f]) - rowoff_1[t]) - coloff_1[f]) - ...      (0 = medians, j >= 1 = iteration j)
// and re-applied in the reference's own order, so r is BIT-IDENTICAL to the reference's given the same offsets.  (Summing the
// offsets first is not an option: the filter divides by ||r|| and a position whose two channels nearly cancel -- they occur,
// ||r|| ~ 1e-9 in the golden case -- amplifies a 1-ulp change of r into 1e-2 of its row / column correction.)
// The passes, all READ-ONLY over x:
//   1. row_median_kernel     warp per row, bitonic sort of the register-resident keys (shuffles)   -> rowoff
//   2. col_hist_kernel x4    column median of (x - rowoff) by 8-bit radix select: shared-memory histograms
//      col_scan_kernel x4    (16-bit counters, two columns per word, 128-bit row loads), warp per column scan
//   3. denoise_iter_kernel x5  warp per row: column offsets of all earlier steps staged in shared memory, norm over the 2
//      col_finalize_kernel x5  channels, num / denom, row means by shuffle, column sums registers -> shared -> fp64 atomics
//   4. denoise_final_kernel  out = r_iters  (the only write)
// Algorithmic traffic 8 B/element (read x, write out); actual 11 reads + 1 write = 48 B/element, all coalesced 128-bit rows.
// The medians are exact (the selected element itself, torch's lower median, rank (n-1)/2); the iterations agree with the
// reference to fp32 rounding (summation order of the means differs).
#include "../../include/ddim_b200.h"
#include "common.cuh"

using namespace ddimb;

namespace {

// monotone map float -> uint32 (ascending), and back
__device__ __forceinline__ uint32_t order_key(float v) {
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(uint32_t k) { return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k); }

// ---------------------------------------------------------------------------------------------- 1. row medians
// Warp-wide bitonic sort of the 32 E keys of one row, E consecutive keys per lane.  Steps whose partner distance is below E
// are register compare-exchanges, the others one shuffle + one predicated min/max per key: ~450 instructions per row against
// ~1600 for a bit-serial radix select (which was compute-bound at 10x the HBM time of the pass).
template <int E>
__device__ __forceinline__ void bitonic_sort_warp(uint32_t (&key)[E], int lane) {
  constexpr int N = 32 * E;
#pragma unroll
  for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= E) {
        const int lj = j / E;
        const bool keepmin = ((lane & lj) == 0) == ((lane & (k / E)) == 0);   // lower index of the pair keeps the min iff ascending
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const uint32_t o = __shfl_xor_sync(0xffffffffu, key[e], lj);
          key[e] = keepmin ? min(key[e], o) : max(key[e], o);
        }
      } else {
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int pe = e ^ j;
          if (pe > e) {
            const bool asc = (k < E) ? ((e & k) == 0) : ((lane & (k / E)) == 0);   // bit k of the global index lane * E + e
            const uint32_t lo = min(key[e], key[pe]), hi = max(key[e], key[pe]);
            key[e] = asc ? lo : hi;
            key[pe] = asc ? hi : lo;
          }
        }
      }
    }
  }
}

template <int FV>   // F = 128 * FV
__global__ void __launch_bounds__(256) row_median_kernel(const float* __restrict__ x, float* __restrict__ rowoff, int64_t rows) {
  constexpr int E = 4 * FV, R = (32 * E - 1) / 2;   // torch.median: the LOWER middle element, rank (n - 1) / 2
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float4* p = reinterpret_cast<const float4*>(x + row * (FV * 128));
  uint32_t key[E];   // any assignment of the row's values to (lane, slot) will do: the sort is over the whole row
#pragma unroll
  for (int i = 0; i < FV; ++i) {
    const float4 v = p[i * 32 + lane];
    key[4 * i + 0] = order_key(v.x);
    key[4 * i + 1] = order_key(v.y);
    key[4 * i + 2] = order_key(v.z);
    key[4 * i + 3] = order_key(v.w);
  }
  bitonic_sort_warp<E>(key, lane);
  if (lane == R / E) rowoff[row] = key_value(key[R % E]);
}

// ---------------------------------------------------------------------------------------------- 2. column medians
// grid (F/128, T chunks, B*C).  A lane owns 4 adjacent columns (one 128-bit load per row), a warp a 512-byte row segment,
// the 8 warps alternate rows.  One radix digit (8 bits at `shift`) of the keys of (x - rowoff) that still match the column's
// prefix is histogrammed in shared memory as 16-bit counters (a chunk has <= 1024 rows), two columns per word:
// sh[bin][64 words], then flushed into hist[col][256] (global, zero on entry).
constexpr int kHistRows = 1024;
constexpr int kHistSmem = 256 * 64 * 4;
__global__ void __launch_bounds__(256) col_hist_kernel(const float* __restrict__ x, const float* __restrict__ rowoff,
                                                       const uint32_t* __restrict__ prefix, uint32_t* __restrict__ hist, int T, int F,
                                                       int shift) {
  extern __shared__ uint32_t sh[];
  for (int i = threadIdx.x; i < 256 * 64; i += 256) sh[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bc = blockIdx.z;
  const int col0 = blockIdx.x * 128 + lane * 4;
  const uint32_t himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
  const uint4 pfx = *reinterpret_cast<const uint4*>(prefix + (int64_t)bc * F + col0);
  const uint32_t pf[4] = {pfx.x, pfx.y, pfx.z, pfx.w};
  const int t0 = blockIdx.y * kHistRows, t1 = min(T, t0 + kHistRows);
  const float* xb = x + (int64_t)bc * T * F + col0;
  const float* rb = rowoff + (int64_t)bc * T;
  for (int t = t0 + warp; t < t1; t += 32) {
    float4 v[4];
    float r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int tt = t + 8 * u;
      v[u] = tt < t1 ? *reinterpret_cast<const float4*>(xb + (int64_t)tt * F) : make_float4(0.f, 0.f, 0.f, 0.f);
      r[u] = tt < t1 ? rb[tt] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (t + 8 * u < t1) {
        const float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t key = order_key(__fsub_rn(e[j], r[u]));
          if ((key & himask) == pf[j]) atomicAdd(&sh[((key >> shift) & 255u) * 64 + lane * 2 + (j >> 1)], (j & 1) ? 0x10000u : 1u);
        }
      }
    }
  }
  __syncthreads();
  uint32_t* hb = hist + ((int64_t)bc * F + blockIdx.x * 128) * 256;
  for (int i = threadIdx.x; i < 256 * 64; i += 256) {
    const uint32_t w = sh[i];
    if (w) {
      const int c = (i & 63) * 2, bin = i >> 6;
      if (w & 0xffffu) atomicAdd(hb + c * 256 + bin, w & 0xffffu);
      if (w >> 16) atomicAdd(hb + (c + 1) * 256 + bin, w >> 16);
    }
  }
}

// warp per column: find the digit holding rank krem, extend the prefix, re-zero the histogram for the next digit.
__global__ void __launch_bounds__(256) col_scan_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ prefix, uint32_t* __restrict__ krem,
                                                       float* __restrict__ coloff, int64_t ncols, int shift) {
  const int lane = threadIdx.x & 31;
  const int64_t col = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (col >= ncols) return;
  uint4* h = reinterpret_cast<uint4*>(hist + col * 256) + lane * 2;
  const uint4 a = h[0], b = h[1];
  h[0] = make_uint4(0, 0, 0, 0);
  h[1] = make_uint4(0, 0, 0, 0);
  const uint32_t c[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += c[j];
  uint32_t incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  const uint32_t excl = incl - s;
  const uint32_t k = krem[col];
  __syncwarp();
  if (k >= excl && k < incl) {
    uint32_t run = excl, digit = 0, knew = 0;
    bool found = false;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (!found && k < run + c[j]) {
        digit = lane * 8 + j;
        knew = k - run;
        found = true;
      }
      run += c[j];
    }
    const uint32_t p = prefix[col] | (digit << shift);
    prefix[col] = p;
    krem[col] = knew;
    if (shift == 0) coloff[col] = key_value(p);
  }
}

__global__ void __launch_bounds__(256) col_select_init_kernel(uint32_t* __restrict__ prefix, uint32_t* __restrict__ krem, int64_t ncols, uint32_t k0) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < ncols) {
    prefix[i] = 0;
    krem[i] = k0;
  }
}

// ---------------------------------------------------------------------------------------------- 3. the five iterations
// grid (T chunks, B).  Warp per row t (both channels): r = x minus the offsets of steps 0..K-1; denom = 1 / (1e-9 + sqrt(r0^2 + r1^2));
// num_c = r_c denom.  Row correction of step K (denoise.py:66-69, axis -1) = mean_f(num_c) / mean_f(denom) -> rowoff[K];
// the column sums of num_0, num_1, denom go to colsum[b][3][F] (double).
template <int FV>
__global__ void __launch_bounds__(256) denoise_iter_kernel(const float* __restrict__ x, float* __restrict__ rowoff, const float* __restrict__ coloff,
                                                           double* __restrict__ colsum, int T, int rpc, int K, int64_t row_stride,
                                                           int64_t col_stride) {
  constexpr int F = FV * 128;
  extern __shared__ float dsm[];
  float* red = dsm;                 // [3][F]
  float* cof = dsm + 3 * F;         // [K][2][F]: column offsets of steps 0..K-1 of this sample
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < K * 2 * F; i += 256) {
    const int j = i / (2 * F), rem = i - j * 2 * F;
    cof[i] = coloff[j * col_stride + (int64_t)(2 * b) * F + rem];
  }
  __syncthreads();
  float an0[4 * FV], an1[4 * FV], ad[4 * FV];
#pragma unroll
  for (int j = 0; j < 4 * FV; ++j) an0[j] = an1[j] = ad[j] = 0.f;
  const int t1 = min(T, (int)blockIdx.x * rpc + rpc);
  float* ro0p = rowoff + (int64_t)(2 * b) * T;
  float* ro1p = rowoff + (int64_t)(2 * b + 1) * T;
  const float* x0p = x + (int64_t)(2 * b) * T * F;
  const float* x1p = x + (int64_t)(2 * b + 1) * T * F;
  constexpr float invF = 1.0f / F;
  for (int t = blockIdx.x * rpc + warp; t < t1; t += 8) {
    float4 v0[FV], v1[FV];
#pragma unroll
    for (int i = 0; i < FV; ++i) {
      v0[i] = reinterpret_cast<const float4*>(x0p + (int64_t)t * F)[i * 32 + lane];
      v1[i] = reinterpret_cast<const float4*>(x1p + (int64_t)t * F)[i * 32 + lane];
    }
    for (int j = 0; j < K; ++j) {   // the reference's order: row offset of step j, then its column offset
      const float ro0 = ro0p[j * row_stride + t], ro1 = ro1p[j * row_stride + t];
#pragma unroll
      for (int i = 0; i < FV; ++i) {
        const float4 o0 = reinterpret_cast<const float4*>(cof + (j * 2) * F)[i * 32 + lane];
        const float4 o1 = reinterpret_cast<const float4*>(cof + (j * 2 + 1) * F)[i * 32 + lane];
        v0[i].x = __fsub_rn(__fsub_rn(v0[i].x, ro0), o0.x); v0[i].y = __fsub_rn(__fsub_rn(v0[i].y, ro0), o0.y);
        v0[i].z = __fsub_rn(__fsub_rn(v0[i].z, ro0), o0.z); v0[i].w = __fsub_rn(__fsub_rn(v0[i].w, ro0), o0.w);
        v1[i].x = __fsub_rn(__fsub_rn(v1[i].x, ro1), o1.x); v1[i].y = __fsub_rn(__fsub_rn(v1[i].y, ro1), o1.y);
        v1[i].z = __fsub_rn(__fsub_rn(v1[i].z, ro1), o1.z); v1[i].w = __fsub_rn(__fsub_rn(v1[i].w, ro1), o1.w);
      }
    }
    float sn0 = 0.f, sn1 = 0.f, sd = 0.f;
#pragma unroll
    for (int i = 0; i < FV; ++i) {
      const float a[4] = {v0[i].x, v0[i].y, v0[i].z, v0[i].w};
      const float c[4] = {v1[i].x, v1[i].y, v1[i].z, v1[i].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float r0 = a[j], r1 = c[j];
        const float d = __frcp_rn(1e-9f + __fsqrt_rn(__fadd_rn(__fmul_rn(r0, r0), __fmul_rn(r1, r1))));
        const float n0 = r0 * d, n1 = r1 * d;
        sn0 += n0; sn1 += n1; sd += d;
        an0[4 * i + j] += n0; an1[4 * i + j] += n1; ad[4 * i + j] += d;
      }
    }
    sn0 = warp_sum(sn0);
    sn1 = warp_sum(sn1);
    sd = warp_sum(sd);
    if (lane == 0) {               // row correction of step K (denoise.py:66-69, axis -1)
      const float md = sd * invF;
      ro0p[K * row_stride + t] = __fdiv_rn(sn0 * invF, md);
      ro1p[K * row_stride + t] = __fdiv_rn(sn1 * invF, md);
    }
  }
  // deterministic in-CTA reduction of the column sums: warps add in turn
  for (int w = 0; w < 8; ++w) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < FV; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int f = i * 128 + lane * 4 + j;
          if (w == 0) {
            red[f] = an0[4 * i + j]; red[F + f] = an1[4 * i + j]; red[2 * F + f] = ad[4 * i + j];
          } else {
            red[f] += an0[4 * i + j]; red[F + f] += an1[4 * i + j]; red[2 * F + f] += ad[4 * i + j];
          }
        }
      }
    }
    __syncthreads();
  }
  double* cs = colsum + (int64_t)b * 3 * F;
  for (int i = threadIdx.x; i < 3 * F; i += 256) atomicAdd(cs + i, (double)red[i]);
}

// column correction of this step (denoise.py:66-69, axis -2) = mean_t(num_c) / mean_t(denom); colsum re-zeroed for the next pass.
__global__ void __launch_bounds__(256) col_finalize_kernel(double* __restrict__ colsum, float* __restrict__ coloff, int B, int F, int T) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= B * F) return;
  const int b = i / F, f = i - b * F;
  double* cs = colsum + (int64_t)b * 3 * F;
  const float md = (float)(cs[2 * F + f] / T);
  const float m0 = (float)(cs[f] / T), m1 = (float)(cs[F + f] / T);
  cs[f] = 0.0; cs[F + f] = 0.0; cs[2 * F + f] = 0.0;
  coloff[(int64_t)(2 * b) * F + f] = __fdiv_rn(m0, md);
  coloff[(int64_t)(2 * b + 1) * F + f] = __fdiv_rn(m1, md);
}

// ---------------------------------------------------------------------------------------------- 4. output
__global__ void __launch_bounds__(256) denoise_final_kernel(const float4* __restrict__ x, const float* __restrict__ rowoff,
                                                            const float* __restrict__ coloff, float4* __restrict__ out, int T, int F4,
                                                            int64_t n4, int K, int64_t row_stride, int64_t col_stride) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / F4;
    const int f4 = (int)(i - row * F4);
    const int64_t bc = row / T;
    float4 v = x[i];
    for (int j = 0; j < K; ++j) {
      const float ro = rowoff[j * row_stride + row];
      const float4 co = reinterpret_cast<const float4*>(coloff + j * col_stride + bc * (F4 * 4))[f4];
      v.x = __fsub_rn(__fsub_rn(v.x, ro), co.x);
      v.y = __fsub_rn(__fsub_rn(v.y, ro), co.y);
      v.z = __fsub_rn(__fsub_rn(v.z, ro), co.z);
      v.w = __fsub_rn(__fsub_rn(v.w, ro), co.w);
    }
    out[i] = v;
  }
}

struct Workspace {
  float* rowoff;
  float* coloff;
  uint32_t* prefix;
  uint32_t* krem;
  double* colsum;
  uint32_t* hist;
  size_t hist_bytes, colsum_bytes, total;
};

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

Workspace carve(void* base, int B, int C, int T, int F, int iters) {
  Workspace w;
  const size_t bc = (size_t)B * C;
  size_t off = 0;
  char* p = static_cast<char*>(base);
  w.rowoff = reinterpret_cast<float*>(p + off);   off += align256((size_t)(iters + 1) * bc * T * sizeof(float));   // [iters+1][B*C][T]
  w.coloff = reinterpret_cast<float*>(p + off);   off += align256((size_t)(iters + 1) * bc * F * sizeof(float));   // [iters+1][B*C][F]
  w.prefix = reinterpret_cast<uint32_t*>(p + off); off += align256(bc * F * sizeof(uint32_t));
  w.krem = reinterpret_cast<uint32_t*>(p + off);  off += align256(bc * F * sizeof(uint32_t));
  w.colsum_bytes = (size_t)B * 3 * F * sizeof(double);
  w.colsum = reinterpret_cast<double*>(p + off);  off += align256(w.colsum_bytes);
  w.hist_bytes = bc * F * 256 * sizeof(uint32_t);
  w.hist = reinterpret_cast<uint32_t*>(p + off);  off += align256(w.hist_bytes);
  w.total = off;
  return w;
}

template <int FV>
int run(const float* x, float* out, const Workspace& w, int B, int T, int iters, cudaStream_t stream) {
  constexpr int F = FV * 128, C = 2;
  const int64_t rows = (int64_t)B * C * T, ncols = (int64_t)B * C * F;
  row_median_kernel<FV><<<(unsigned)((rows + 7) / 8), 256, 0, stream>>>(x, w.rowoff, rows);
  DDIMB_LAUNCH_CHECK();
  DDIMB_CUDA(cudaMemsetAsync(w.hist, 0, w.hist_bytes, stream));
  DDIMB_CUDA(cudaMemsetAsync(w.colsum, 0, w.colsum_bytes, stream));
  col_select_init_kernel<<<(unsigned)((ncols + 255) / 256), 256, 0, stream>>>(w.prefix, w.krem, ncols, (uint32_t)((T - 1) / 2));
  DDIMB_LAUNCH_CHECK();
  DDIMB_SMEM_ATTR(col_hist_kernel, kHistSmem);   // once per device
  for (int shift = 24; shift >= 0; shift -= 8) {
    col_hist_kernel<<<dim3(F / 128, (T + kHistRows - 1) / kHistRows, B * C), 256, kHistSmem, stream>>>(x, w.rowoff, w.prefix, w.hist, T, F, shift);
    DDIMB_LAUNCH_CHECK();
    col_scan_kernel<<<(unsigned)((ncols + 7) / 8), 256, 0, stream>>>(w.hist, w.prefix, w.krem, w.coloff, ncols, shift);
    DDIMB_LAUNCH_CHECK();
  }
  // rows per CTA of the iteration pass: enough CTAs for ~4 per SM, at most 256 rows (column accumulators stay fp32-accurate)
  const int want = (4 * ddimb_num_sms() + B - 1) / B;
  int rpc = ((T + want - 1) / want + 7) / 8 * 8;
  rpc = rpc < 8 ? 8 : (rpc > 256 ? 256 : rpc);
  for (int it = 0; it < iters; ++it) {
    const int K = it + 1;   // offsets of steps 0..it are applied, step K is produced
    const size_t smem = (size_t)(3 + 2 * K) * F * sizeof(float);
    denoise_iter_kernel<FV><<<dim3((T + rpc - 1) / rpc, B), 256, smem, stream>>>(x, w.rowoff, w.coloff, w.colsum, T, rpc, K, rows, ncols);
    DDIMB_LAUNCH_CHECK();
    col_finalize_kernel<<<(B * F + 255) / 256, 256, 0, stream>>>(w.colsum, w.coloff + (int64_t)K * ncols, B, F, T);
    DDIMB_LAUNCH_CHECK();
  }
  const int64_t n4 = rows * (F / 4);
  int64_t blocks = (n4 + 255) / 256;
  const int64_t cap = (int64_t)ddimb_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  denoise_final_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(x), w.rowoff, w.coloff, reinterpret_cast<float4*>(out),
                                                              T, F / 4, n4, iters + 1, rows, ncols);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

}  // namespace

extern "C" long long ddimb_denoise2d_workspace(int B, int C, int T, int F, int iters) {
  if (B < 1 || C != 2 || T < 1 || (F != 128 && F != 256 && F != 512) || iters < 0 || iters > 8) return -1;
  return (long long)carve(nullptr, B, C, T, F, iters).total;
}

extern "C" int ddimb_denoise2d(const float* x, float* out, void* workspace, long long workspace_bytes, int B, int C, int T, int F, int iters,
                               cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && out && workspace && B >= 1 && T >= 1 && iters >= 0 && iters <= 8);   // (3 + 2 iters) F floats of shared memory
  DDIMB_CHECK_ARG(C == 2);                                     // audio.yml channels; the norm of denoise.py:64 runs over this axis
  DDIMB_CHECK_ARG(F == 128 || F == 256 || F == 512);
  DDIMB_CHECK_ARG((long long)B * C <= 65535);                  // grid.z / grid.y limits
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(workspace) & 255) == 0);
  const Workspace w = carve(workspace, B, C, T, F, iters);
  DDIMB_CHECK_ARG(workspace_bytes >= (long long)w.total);
  DDIMB_PROF(stream, "denoise2d", F, iters, 8.0 * B * C * T * F, 12.0 * iters * B * C * T * F);
  if (F == 128) return run<1>(x, out, w, B, T, iters, stream);
  if (F == 256) return run<2>(x, out, w, B, T, iters, stream);
  return run<4>(x, out, w, B, T, iters, stream);
}

// ------------------------------------------------------------------------------------------------ allocation-free variant
// No counterpart in the reference: added for callers that cannot allocate device memory (code
// under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Same kernels as ddimb_denoise2d, but the scratch is the
// file-scope device array below instead of a caller buffer: every GPU that loads the library holds kDenoiseStaticBytes until exit.
// The array is shared by every call on that GPU and keeps the previous call's contents -- the parts that must start at zero
// (histograms, column sums) are cleared on the stream at the start of every call, exactly as the caller-workspace path does -- so
// two calls that overlap on the GPU (different streams, no ordering) race on it.
constexpr long long kDenoiseStaticBytes = 4ll << 20;           // 4 MiB: e.g. x[1,2,8192,256] needs 1.06 MiB, x[3,2,8192,256] 3.2 MiB at 5 iterations
static __device__ __align__(256) unsigned char g_denoise_scratch[kDenoiseStaticBytes];

extern "C" long long ddimb_denoise2d_static_max_bytes() { return kDenoiseStaticBytes; }

extern "C" int ddimb_denoise2d_static(const float* x, float* out, int B, int C, int T, int F, int iters, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && out && B >= 1 && T >= 1 && iters >= 0 && iters <= 8);
  DDIMB_CHECK_ARG(C == 2);
  DDIMB_CHECK_ARG(F == 128 || F == 256 || F == 512);
  DDIMB_CHECK_ARG((long long)B * C <= 65535);
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0);
  const long long need = (long long)carve(nullptr, B, C, T, F, iters).total;
  if (need > kDenoiseStaticBytes) {
    ddimb_set_error("ddimb_denoise2d_static: x[%d,%d,%d,%d] with %d iterations needs %lld bytes of scratch, the built-in array holds %lld; "
                    "use ddimb_denoise2d with a caller workspace", B, C, T, F, iters, need, kDenoiseStaticBytes);
    return DDIMB_ERR_ARG;
  }
  // In host code the array's name is a host-side stand-in, not a device address: resolve it once per device, with that device current.
  static void* dev_ptr[64] = {nullptr};
  int dev = 0;
  DDIMB_CUDA(cudaGetDevice(&dev));
  DDIMB_CHECK_ARG(dev >= 0 && dev < 64);
  if (!dev_ptr[dev]) {
    void* ptr = nullptr;
    DDIMB_CUDA(cudaGetSymbolAddress(&ptr, g_denoise_scratch));
    DDIMB_CHECK_ARG(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255) == 0);
    dev_ptr[dev] = ptr;
  }
  const Workspace w = carve(dev_ptr[dev], B, C, T, F, iters);
  DDIMB_PROF(stream, "denoise2d_static", F, iters, 8.0 * B * C * T * F, 12.0 * iters * B * C * T * F);
  if (F == 128) return run<1>(x, out, w, B, T, iters, stream);
  if (F == 256) return run<2>(x, out, w, B, T, iters, stream);
  return run<4>(x, out, w, B, T, iters, stream);
}
