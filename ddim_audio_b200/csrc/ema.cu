// EMA of the parameters (reference models/ema.py:16-23, called every step at runners/diffusion.py:182-183) as ONE
// multi-tensor launch -- SURVEY.md section 8 row f3.  The reference walks the ~390 parameters with three torch ops each
// (~1200 launches and two temporaries per tensor per step); here a device table of <= 16 K-element chunks
// {param, shadow, count} is built once at register() and every CTA streams one chunk: 12 B/element (read p, read + write s),
// 128-bit accesses.  Same operation order and roundings as the reference ((1 - mu) p rounded, mu s rounded, then the sum):
// bit-exact.
#include "../../include/ddim_b200.h"
#include "common.cuh"

namespace {
struct EmaChunk {
  const float* p;
  float* s;
  long long n;
  long long pad;
};

__global__ void __launch_bounds__(256) ema_update_kernel(const EmaChunk* __restrict__ table, float one_minus_mu, float mu) {
  const EmaChunk c = table[blockIdx.x];
  const int n4 = ((reinterpret_cast<uintptr_t>(c.p) | reinterpret_cast<uintptr_t>(c.s)) & 15) ? 0 : (int)(c.n >> 2);
  const float4* p4 = reinterpret_cast<const float4*>(c.p);
  float4* s4 = reinterpret_cast<float4*>(c.s);
  for (int i = threadIdx.x; i < n4; i += 256) {
    const float4 p = p4[i];
    float4 s = s4[i];
    s.x = __fadd_rn(__fmul_rn(one_minus_mu, p.x), __fmul_rn(mu, s.x));
    s.y = __fadd_rn(__fmul_rn(one_minus_mu, p.y), __fmul_rn(mu, s.y));
    s.z = __fadd_rn(__fmul_rn(one_minus_mu, p.z), __fmul_rn(mu, s.z));
    s.w = __fadd_rn(__fmul_rn(one_minus_mu, p.w), __fmul_rn(mu, s.w));
    s4[i] = s;
  }
  for (long long i = (long long)n4 * 4 + threadIdx.x; i < c.n; i += 256)
    c.s[i] = __fadd_rn(__fmul_rn(one_minus_mu, c.p[i]), __fmul_rn(mu, c.s[i]));
}
}  // namespace

extern "C" int ddimb_ema_update(const void* chunk_table, int nchunks, long long total_elems, double mu, cudaStream_t stream) {
  DDIMB_CHECK_ARG(chunk_table && nchunks >= 1 && total_elems >= 1);
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(chunk_table) & 15) == 0);
  DDIMB_PROF(stream, "ema_update", 0, 0, 12.0 * total_elems, 3.0 * total_elems);
  // torch multiplies a float tensor by a Python scalar in fp32: both factors are rounded to float first
  ema_update_kernel<<<nchunks, 256, 0, stream>>>(static_cast<const EmaChunk*>(chunk_table), (float)(1.0 - mu), (float)mu);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ weight re-pack (training)
// The GEMM kernels read bf16 K-major operand copies of the fp32 master weights in several layouts (models/diffusion.py:28-78 weights as
// include/ddim_b200.h lays them out: 3x3 kernels tap-major, their flipped + transposed data-gradient twins, 4x4 kernels, Linear weights
// and their transposes).  During training the masters change every step; rebuilding the copies through ~600 torch permute / flip /
// cast / contiguous calls cost 25 ms of HOST time per step (the train step was host-bound).  Every such layout is a strided gather
//     dst[i0][i1][i2][i3] = src[off + i0 s0 + i1 s1 + i2 s2 + i3 s3]      (negative strides = flips)
// so ONE launch over a chunk table rewrites all copies in place.  A chunk is self-contained: {src, dst, first element, count, inner
// dims, strides}; one CTA per chunk.
struct RepackChunk {
  const float* src;        // fp32 master (element offset `off` already applied)
  void* dst;               // bf16 (or fp32 when dst_f32) copy, contiguous [d0][d1][d2][d3]
  long long start, n;      // this chunk: dst elements [start, start + n)
  int d1, d2, d3, dst_f32;
  long long s0, s1, s2, s3;
};
static_assert(sizeof(RepackChunk) == 80, "RepackChunk layout is part of the ABI");

__global__ void __launch_bounds__(256) repack_kernel(const RepackChunk* __restrict__ table) {
  const RepackChunk c = table[blockIdx.x];
  const long long d23 = (long long)c.d2 * c.d3, d123 = d23 * c.d1;
  for (long long k = threadIdx.x; k < c.n; k += blockDim.x) {
    const long long i = c.start + k;
    const long long i0 = i / d123, r0 = i - i0 * d123;
    const long long i1 = r0 / d23, r1 = r0 - i1 * d23;
    const long long i2 = r1 / c.d3, i3 = r1 - i2 * c.d3;
    const float v = __ldg(c.src + i0 * c.s0 + i1 * c.s1 + i2 * c.s2 + i3 * c.s3);
    if (c.dst_f32) static_cast<float*>(c.dst)[i] = v;
    else static_cast<__nv_bfloat16*>(c.dst)[i] = __float2bfloat16_rn(v);
  }
}

extern "C" int ddimb_repack(const void* chunk_table, int nchunks, long long total_elems, cudaStream_t stream) {
  DDIMB_CHECK_ARG(chunk_table && nchunks >= 1 && total_elems >= 1);
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(chunk_table) & 15) == 0);
  DDIMB_PROF(stream, "repack", 0, 0, 6.0 * total_elems, 0.0);
  repack_kernel<<<nchunks, 256, 0, stream>>>(static_cast<const RepackChunk*>(chunk_table));
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}
