// Optimizer tail of the train step as multi-tensor launches -- SURVEY.md section 8 row f2.
//   reference: runners/diffusion.py:155-173 -- per group torch.nn.utils.clip_grad_norm_, then optimizer.step():
//   AdaBelief (External/step-clip-optimizer/clip_opt/adabelief/AdaBelief.py:55-162; default group, ~330 tensors, ~12 torch ops
//   per tensor from a Python loop) and torch.optim.AdamW (transformer group).
// Here every (group, operation) is ONE launch over a device table of <= 64 K-element chunks
//   {param, grad, exp_avg, exp_avg_var|exp_avg_sq, count}: gradient sum of squares (fp64), in-place clip scaling, and the
// update itself -- 28 B/element (read p, g, m, v; write p, m, v), 128-bit accesses, fp32 math in the reference's operation
// order (scalars formed in double on the host as Python does, then rounded to fp32 where torch rounds them).
#include "../../include/ddim_b200.h"
#include "common.cuh"

using namespace ddimb;

namespace {
struct OptChunk {
  float* p;
  float* g;
  float* m;
  float* v;
  long long n;
  long long pad[3];
};
static_assert(sizeof(OptChunk) == 64, "chunk record layout is part of the ABI");

// 128-bit path only when every pointer of the record allows it (views into foreign flat buffers may sit on any 4-byte offset)
__device__ __forceinline__ int vec4_count(const OptChunk& c) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(c.p) | reinterpret_cast<uintptr_t>(c.g) | reinterpret_cast<uintptr_t>(c.m) |
                      reinterpret_cast<uintptr_t>(c.v);
  return (a & 15) ? 0 : (int)(c.n >> 2);
}

__global__ void __launch_bounds__(256) grad_sqnorm_kernel(const OptChunk* __restrict__ table, double* __restrict__ out) {
  const OptChunk c = table[blockIdx.x];
  const int n4 = vec4_count(c);
  const float4* g4 = reinterpret_cast<const float4*>(c.g);
  float acc = 0.f;
  for (int i = threadIdx.x; i < n4; i += 256) {
    const float4 g = g4[i];
    acc += g.x * g.x + g.y * g.y + g.z * g.z + g.w * g.w;
  }
  for (long long i = (long long)n4 * 4 + threadIdx.x; i < c.n; i += 256) acc += c.g[i] * c.g[i];
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(out, (double)v);
  }
}

// torch.nn.utils.clip_grad_norm_: coef = min(1, max_norm / (total_norm + 1e-6)) in fp32; grads *= coef.  norm_out = total_norm.
__global__ void __launch_bounds__(256) grad_clip_kernel(const OptChunk* __restrict__ table, const double* __restrict__ sqnorm, float max_norm,
                                                        float* __restrict__ norm_out) {
  const float total = (float)sqrt(*sqnorm);
  if (blockIdx.x == 0 && threadIdx.x == 0 && norm_out) *norm_out = total;
  // torch clamps with max=1.0, which lets a NaN total norm through to every gradient (the divergence shows up loudly); fminf would
  // silently return 1
  const float c0 = __fdiv_rn(max_norm, total + 1e-6f);
  const float coef = (c0 < 1.0f || c0 != c0) ? c0 : 1.0f;
  const OptChunk c = table[blockIdx.x];
  const int n4 = vec4_count(c);
  float4* g4 = reinterpret_cast<float4*>(c.g);
  for (int i = threadIdx.x; i < n4; i += 256) {
    float4 g = g4[i];
    g.x *= coef; g.y *= coef; g.z *= coef; g.w *= coef;
    g4[i] = g;
  }
  for (long long i = (long long)n4 * 4 + threadIdx.x; i < c.n; i += 256) c.g[i] *= coef;
}

struct StepScalars {
  float beta1, one_minus_beta1, beta2, one_minus_beta2, bc2_sqrt, eps, decay_mul, neg_step_size;
};

// KIND 0: torch.optim.AdamW (decoupled decay p *= 1 - lr wd; exp_avg.lerp_(g, 1 - b1); exp_avg_sq = b2 v + (1 - b2) g g)
// KIND 1: AdaBelief.py:112-142 (exp_avg = b1 m + (1 - b1) g; residual = g - exp_avg; exp_avg_var = b2 v + (1 - b2) r r;
//         p *= 1 - wd (weight_decouple, not fixed_decay); no amsgrad, no clip_step)
// both: denom = sqrt(v) / sqrt(bias_correction2) + eps;  p += (-lr / bias_correction1) * m / denom
template <int KIND>
__device__ __forceinline__ void update1(float& p, float g, float& m, float& v, const StepScalars& s) {
  float r;
  if (KIND == 0) {
    p = __fmul_rn(p, s.decay_mul);
    m = __fadd_rn(m, __fmul_rn(s.one_minus_beta1, __fsub_rn(g, m)));
    r = g;
  } else {
    m = __fadd_rn(__fmul_rn(m, s.beta1), __fmul_rn(s.one_minus_beta1, g));
    r = __fsub_rn(g, m);
  }
  v = __fadd_rn(__fmul_rn(v, s.beta2), __fmul_rn(__fmul_rn(s.one_minus_beta2, r), r));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), s.bc2_sqrt), s.eps);
  if (KIND == 1) p = __fmul_rn(p, s.decay_mul);
  p = __fadd_rn(p, __fmul_rn(s.neg_step_size, __fdiv_rn(m, denom)));
}

template <int KIND>
__global__ void __launch_bounds__(256) optim_step_kernel(const OptChunk* __restrict__ table, StepScalars s) {
  const OptChunk c = table[blockIdx.x];
  const int n4 = vec4_count(c);
  float4* p4 = reinterpret_cast<float4*>(c.p);
  const float4* g4 = reinterpret_cast<const float4*>(c.g);
  float4* m4 = reinterpret_cast<float4*>(c.m);
  float4* v4 = reinterpret_cast<float4*>(c.v);
  for (int i = threadIdx.x; i < n4; i += 256) {
    float4 p = p4[i], m = m4[i], v = v4[i];
    const float4 g = g4[i];
    update1<KIND>(p.x, g.x, m.x, v.x, s);
    update1<KIND>(p.y, g.y, m.y, v.y, s);
    update1<KIND>(p.z, g.z, m.z, v.z, s);
    update1<KIND>(p.w, g.w, m.w, v.w, s);
    p4[i] = p; m4[i] = m; v4[i] = v;
  }
  for (long long i = (long long)n4 * 4 + threadIdx.x; i < c.n; i += 256) {
    float p = c.p[i], m = c.m[i], v = c.v[i];
    update1<KIND>(p, c.g[i], m, v, s);
    c.p[i] = p; c.m[i] = m; c.v[i] = v;
  }
}
}  // namespace

extern "C" int ddimb_grad_clip(const void* chunk_table, int nchunks, long long total_elems, double max_norm, double* sqnorm_scratch,
                               float* total_norm_out, cudaStream_t stream) {
  DDIMB_CHECK_ARG(chunk_table && nchunks >= 1 && sqnorm_scratch && max_norm > 0.0);
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(chunk_table) & 15) == 0);
  DDIMB_PROF(stream, "grad_clip", 0, 0, 12.0 * total_elems, 3.0 * total_elems);
  DDIMB_CUDA(cudaMemsetAsync(sqnorm_scratch, 0, sizeof(double), stream));
  grad_sqnorm_kernel<<<nchunks, 256, 0, stream>>>(static_cast<const OptChunk*>(chunk_table), sqnorm_scratch);
  DDIMB_LAUNCH_CHECK();
  grad_clip_kernel<<<nchunks, 256, 0, stream>>>(static_cast<const OptChunk*>(chunk_table), sqnorm_scratch, (float)max_norm, total_norm_out);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

extern "C" int ddimb_optim_step(const void* chunk_table, int nchunks, long long total_elems, int kind, double lr, double beta1, double beta2,
                                double eps, double weight_decay, long long step, cudaStream_t stream) {
  DDIMB_CHECK_ARG(chunk_table && nchunks >= 1 && (kind == 0 || kind == 1) && step >= 1);
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(chunk_table) & 15) == 0);
  DDIMB_PROF(stream, "optim_step", kind, 0, 28.0 * total_elems, 14.0 * total_elems);
  // the scalars as Python / torch form them: float64 on the host, rounded to fp32 at the tensor op
  const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
  StepScalars s;
  s.beta1 = (float)beta1;
  s.one_minus_beta1 = (float)(1.0 - beta1);
  s.beta2 = (float)beta2;
  s.one_minus_beta2 = (float)(1.0 - beta2);
  s.bc2_sqrt = (float)sqrt(bc2);
  s.eps = (float)eps;
  s.decay_mul = (float)(kind == 0 ? 1.0 - lr * weight_decay : 1.0 - weight_decay);
  s.neg_step_size = (float)(-lr / bc1);
  if (kind == 0)
    optim_step_kernel<0><<<nchunks, 256, 0, stream>>>(static_cast<const OptChunk*>(chunk_table), s);
  else
    optim_step_kernel<1><<<nchunks, 256, 0, stream>>>(static_cast<const OptChunk*>(chunk_table), s);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}
