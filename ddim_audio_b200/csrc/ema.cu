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
