// HBM-bound kernels of the denoiser path: 128-bit vectorised, fp32 math, bf16 NHWC storage.
// Every GroupNorm statistic is a pair (sum, sum of squares) per (sample, group) accumulated in double by the
// kernel that PRODUCES the tensor (conv epilogue or one of the kernels below), so no kernel re-reads a tensor
// just to normalise it.
#include "../../include/ddim_b200.h"
#include "common.cuh"
#include <stdlib.h>

using namespace ddimb;

// ------------------------------------------------------------------------------------------------ DDIM update
// reference functions/denoising.py:27,41-43.  In place on xt (fp32), same operation order as the reference:
//   x0 = (xt + (-sqrt(1-at)) * et) / sqrt(at);   xt' = x0 * sqrt(at_next) + c2 * et + c1 * z
__global__ void __launch_bounds__(256) ddim_step_kernel(float4* __restrict__ xt, const float4* __restrict__ et,
                                                        const float4* __restrict__ z, float4* __restrict__ x0_out,
                                                        float4* __restrict__ xs_out, float neg_s1mat, float sat, float sat_next,
                                                        float c2, float c1, int64_t n4) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 x = xt[i];
    const float4 e = et[i];
    x.x = __fdiv_rn(fmaf(neg_s1mat, e.x, x.x), sat);
    x.y = __fdiv_rn(fmaf(neg_s1mat, e.y, x.y), sat);
    x.z = __fdiv_rn(fmaf(neg_s1mat, e.z, x.z), sat);
    x.w = __fdiv_rn(fmaf(neg_s1mat, e.w, x.w), sat);
    if (x0_out) x0_out[i] = x;
    x.x = fmaf(c2, e.x, x.x * sat_next);
    x.y = fmaf(c2, e.y, x.y * sat_next);
    x.z = fmaf(c2, e.z, x.z * sat_next);
    x.w = fmaf(c2, e.w, x.w * sat_next);
    if (z) {
      const float4 n = z[i];
      x.x = fmaf(c1, n.x, x.x); x.y = fmaf(c1, n.y, x.y); x.z = fmaf(c1, n.z, x.z); x.w = fmaf(c1, n.w, x.w);
    }
    xt[i] = x;
    if (xs_out) xs_out[i] = x;
  }
}

extern "C" int ddimb_ddim_step(float* xt, const float* et, const float* z, float* x0_out, float* xs_out, double at, double at_next,
                               double c1, double c2, long long n, cudaStream_t stream) {
  DDIMB_CHECK_ARG(xt && et && n > 0 && n % 4 == 0);
  DDIMB_PROF(stream, "ddim_step", (z && c1 != 0.0) ? 1 : 0, (x0_out ? 1 : 0) + (xs_out ? 1 : 0),
             4.0 * n * (3 + ((z && c1 != 0.0) ? 1 : 0) + (x0_out ? 1 : 0) + (xs_out ? 1 : 0)), 6.0 * n);
  const int64_t n4 = n / 4;
  int64_t blocks = (n4 + 255) / 256;
  const int64_t cap = (int64_t)ddimb_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  // the reference forms the scalars in float64 on the host and torch casts them to fp32 at the op boundary
  ddim_step_kernel<<<(int)blocks, 256, 0, stream>>>(reinterpret_cast<float4*>(xt), reinterpret_cast<const float4*>(et),
                                                    (c1 != 0.0 && z) ? reinterpret_cast<const float4*>(z) : nullptr,
                                                    reinterpret_cast<float4*>(x0_out), reinterpret_cast<float4*>(xs_out),
                                                    (float)(-sqrt(1.0 - at)), (float)sqrt(at), (float)sqrt(at_next), (float)c2,
                                                    (float)c1, n4);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ q-sample / loss
// reference functions/losses.py:12-13: x = x0 * sqrt(a[t]) + e * sqrt(1 - a[t])
__global__ void __launch_bounds__(256) qsample_kernel(const float4* __restrict__ x0, const float4* __restrict__ e,
                                                      const float* __restrict__ alphas, const int64_t* __restrict__ t,
                                                      float4* __restrict__ out, int64_t per4) {
  const int b = blockIdx.y;
  const float a = alphas[t[b]];
  const float sa = sqrtf(a), sb = sqrtf(1.0f - a);
  const int64_t base = (int64_t)b * per4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 x = x0[base + i], n = e[base + i];
    out[base + i] = make_float4(fmaf(n.x, sb, x.x * sa), fmaf(n.y, sb, x.y * sa), fmaf(n.z, sb, x.z * sa), fmaf(n.w, sb, x.w * sa));
  }
}

extern "C" int ddimb_qsample(const float* x0, const float* e, const float* alphas, const long long* t, float* out, int B,
                             long long per_sample, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x0 && e && alphas && t && out && B >= 1 && per_sample % 4 == 0);
  DDIMB_PROF(stream, "qsample", 0, 0, 12.0 * B * per_sample, 3.0 * B * per_sample);
  const int64_t per4 = per_sample / 4;
  int bx = (int)((per4 + 255) / 256);
  const int cap = ddimb_num_sms() * 8 / B + 1;
  if (bx > cap) bx = cap;
  qsample_kernel<<<dim3(bx, B), 256, 0, stream>>>(reinterpret_cast<const float4*>(x0), reinterpret_cast<const float4*>(e), alphas,
                                                  reinterpret_cast<const int64_t*>(t), reinterpret_cast<float4*>(out), per4);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// reference functions/losses.py:15-18: per-sample sum over (C,T,F) of (e - output)^2, accumulated in double.
__global__ void __launch_bounds__(256) sqerr_kernel(const float4* __restrict__ e, const float4* __restrict__ o,
                                                    double* __restrict__ per_sample, int64_t per4) {
  const int b = blockIdx.y;
  const int64_t base = (int64_t)b * per4;
  float acc = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = e[base + i], c = o[base + i];
    const float d0 = a.x - c.x, d1 = a.y - c.y, d2 = a.z - c.z, d3 = a.w - c.w;
    acc += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(per_sample + b, (double)v);
  }
}

extern "C" int ddimb_sqerr_loss(const float* e, const float* output, double* per_sample, int B, long long per_sample_elems,
                                cudaStream_t stream) {
  DDIMB_CHECK_ARG(e && output && per_sample && B >= 1 && per_sample_elems % 4 == 0);
  DDIMB_PROF(stream, "sqerr_loss", 0, 0, 8.0 * B * per_sample_elems, 3.0 * B * per_sample_elems);
  DDIMB_CUDA(cudaMemsetAsync(per_sample, 0, sizeof(double) * B, stream));
  const int64_t per4 = per_sample_elems / 4;
  int bx = (int)((per4 + 255) / 256);
  const int cap = ddimb_num_sms() * 8 / B + 1;
  if (bx > cap) bx = cap;
  sqerr_kernel<<<dim3(bx, B), 256, 0, stream>>>(reinterpret_cast<const float4*>(e), reinterpret_cast<const float4*>(output), per_sample, per4);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// backward of the reduction above: grad_output = -2 (e - output) * g[b]
__global__ void __launch_bounds__(256) sqerr_bwd_kernel(const float4* __restrict__ e, const float4* __restrict__ o,
                                                        const float* __restrict__ g, float4* __restrict__ go, int64_t per4) {
  const int b = blockIdx.y;
  const int64_t base = (int64_t)b * per4;
  const float s = -2.0f * g[b];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = e[base + i], c = o[base + i];
    go[base + i] = make_float4(s * (a.x - c.x), s * (a.y - c.y), s * (a.z - c.z), s * (a.w - c.w));
  }
}

extern "C" int ddimb_sqerr_loss_bwd(const float* e, const float* output, const float* g, float* grad_output, int B,
                                    long long per_sample_elems, cudaStream_t stream) {
  DDIMB_CHECK_ARG(e && output && g && grad_output && B >= 1 && per_sample_elems % 4 == 0);
  DDIMB_PROF(stream, "sqerr_loss_bwd", 0, 0, 12.0 * B * per_sample_elems, 2.0 * B * per_sample_elems);
  const int64_t per4 = per_sample_elems / 4;
  int bx = (int)((per4 + 255) / 256);
  const int cap = ddimb_num_sms() * 8 / B + 1;
  if (bx > cap) bx = cap;
  sqerr_bwd_kernel<<<dim3(bx, B), 256, 0, stream>>>(reinterpret_cast<const float4*>(e), reinterpret_cast<const float4*>(output), g,
                                                    reinterpret_cast<float4*>(grad_output), per4);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ GroupNorm family
// reference models/diffusion.py:17-26 (GroupNorm(8, C, eps=1e-6)), :42-56 (where each one sits in Residual_Block).
// x is NHWC bf16 [B][P][C]; stats[b][g] = (sum, sumsq) over the C/8 channels x P positions of group g.
//   mode 0: y = gn(x)                       (GN1, feeds conv1 directly)
//   mode 1: y = silu(gn(x))                 (GN0 + SiLU, feeds conv0)
//   mode 2: y = res + gn(x), + stats of y   (GN2 + residual add; y is the next block's input)
struct GnArgs {
  const __nv_bfloat16* x;
  const __nv_bfloat16* res;
  const __nv_bfloat16* res2;   // optional second residual (mode 2): y = res + res2 + gn(x) -- the last block also adds conv_out's skip operand
  __nv_bfloat16* y;
  const double* stats_in;
  double* stats_out;
  const float* gamma;
  const float* beta;
  int64_t per_sample;   // P * C
  int C;
  int mode;
  int pre_silu;         // the normalised tensor is silu(x) (training stores pre-activations)
  float eps;
};

// Block reduction of the per-thread GroupNorm partial sums (sa1, sa2: channels c0..c0+3; sb1, sb2: c0+4..c0+7) into the 8 x (sum, sumsq)
// of the sample, then one fp64 atomic per statistic.  Fixed order (run-to-run reproducible per CTA): thread (slot, part) adds the
// partials of threads part, part + 16, ... that belong to group slot / 2 in increasing order, then a fixed shuffle tree joins the 16
// parts.  (The first version had 16 threads walk all 256 partials with a division per step: ~13 us per call, which is most of a
// mode-2 call on the small tensors of the deep levels / small sampling batches.)
__device__ __forceinline__ void gn_block_stats_flush(float (*s_part)[4], unsigned char (*s_grp)[2], float sa1, float sa2, float sb1, float sb2,
                                                     double* stats_out_b) {
  s_part[threadIdx.x][0] = sa1; s_part[threadIdx.x][1] = sa2; s_part[threadIdx.x][2] = sb1; s_part[threadIdx.x][3] = sb2;
  __syncthreads();
  const int part = threadIdx.x & 15;
  // slot = group * 2 + which; a half warp per slot, whole warps per pass (blocks of 192 threads take slots 12..15 in a second pass)
  for (int slot = threadIdx.x >> 4; slot < 16; slot += (int)(blockDim.x >> 4)) {
    const int g = slot >> 1, which = slot & 1;
    float acc = 0.f;
    for (int t = part; t < (int)blockDim.x; t += 16) {
      if (s_grp[t][0] == g) acc += s_part[t][which];
      if (s_grp[t][1] == g) acc += s_part[t][2 + which];
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o, 16);
    if (part == 0) atomicAdd(stats_out_b + slot, (double)acc);
  }
}

// RES2: compile-time presence of the second residual, so the common instantiation keeps its register count / loads in flight
// (as a run-time branch the extra UNR x 16 B registers cost every mode-2 call 10-20 %)
template <int UNR, bool RES2>
__global__ void __launch_bounds__(256) gn_apply_kernel(const GnArgs a) {
  __shared__ float s_scale[256], s_shift[256];
  const int b = blockIdx.y, C = a.C, cg = C >> 3;
  if (threadIdx.x < C) {
    const int c = threadIdx.x, g = c / cg;
    const double n = (double)(a.per_sample / 8);
    const double mean = a.stats_in[(b * 8 + g) * 2] / n;
    double var = a.stats_in[(b * 8 + g) * 2 + 1] / n - mean * mean;
    if (var < 0) var = 0;
    const float rstd = (float)(1.0 / sqrt(var + (double)a.eps));
    const float sc = a.gamma[c] * rstd;
    s_scale[c] = sc;
    s_shift[c] = (a.beta ? a.beta[c] : 0.f) - (float)mean * sc;
  }
  __syncthreads();
  // blockDim * 8 is a multiple of C, so a thread keeps the same 8 channels for every item it visits
  const int c0 = (threadIdx.x * 8) % C;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = s_scale[c0 + j]; sh[j] = s_shift[c0 + j]; }
  const int64_t items = a.per_sample / 8;
  const uint4* xp = reinterpret_cast<const uint4*>(a.x + (int64_t)b * a.per_sample);
  const uint4* rp = a.res ? reinterpret_cast<const uint4*>(a.res + (int64_t)b * a.per_sample) : nullptr;
  const uint4* r2p = a.res2 ? reinterpret_cast<const uint4*>(a.res2 + (int64_t)b * a.per_sample) : nullptr;
  uint4* yp = reinterpret_cast<uint4*>(a.y + (int64_t)b * a.per_sample);
  float sa1 = 0.f, sa2 = 0.f, sb1 = 0.f, sb2 = 0.f;   // channels c0..c0+3 and c0+4..c0+7 (group width is a multiple of 4)
  // UNR items per iteration with every load issued before the first use: ~8 x 16 B in flight per thread keeps HBM busy
  // (UNR = 8 for the one-operand modes, 4 for mode 2 which also loads the residual)
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < items; i0 += UNR * stride) {
    uint4 xv[UNR], rv[UNR], qv[RES2 ? UNR : 1];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < items) {
        xv[u] = ldg_nc_v4(xp + i);
        if (a.mode == 2) rv[u] = ldg_nc_v4(rp + i);
        if (RES2) qv[u] = ldg_nc_v4(r2p + i);
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int64_t i = i0 + u * stride;
      if (i >= items) break;
      const uint32_t xw[4] = {xv[u].x, xv[u].y, xv[u].z, xv[u].w};
      float f[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 t = unpack_bf16x2(xw[j]);
        if (a.pre_silu) { t.x = silu_fast(t.x); t.y = silu_fast(t.y); }
        f[2 * j] = fmaf(t.x, sc[2 * j], sh[2 * j]);
        f[2 * j + 1] = fmaf(t.y, sc[2 * j + 1], sh[2 * j + 1]);
      }
      if (a.mode == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = silu_fast(f[j]);
      } else if (a.mode == 2) {
        const uint32_t rw[4] = {rv[u].x, rv[u].y, rv[u].z, rv[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 t = unpack_bf16x2(rw[j]);
          f[2 * j] += t.x; f[2 * j + 1] += t.y;
        }
        if (RES2) {
          const uint32_t qw[4] = {qv[u].x, qv[u].y, qv[u].z, qv[u].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 t = unpack_bf16x2(qw[j]);
            f[2 * j] += t.x; f[2 * j + 1] += t.y;
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) { sa1 += f[j]; sa2 = fmaf(f[j], f[j], sa2); sb1 += f[4 + j]; sb2 = fmaf(f[4 + j], f[4 + j], sb2); }
      }
      stg_v4(yp + i, make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7])));
    }
  }
  if (a.mode == 2 && a.stats_out) {
    __shared__ float s_part[256][4];
    __shared__ unsigned char s_grp[256][2];
    s_grp[threadIdx.x][0] = (unsigned char)(c0 / cg); s_grp[threadIdx.x][1] = (unsigned char)((c0 + 4) / cg);
    gn_block_stats_flush(s_part, s_grp, sa1, sa2, sb1, sb2, a.stats_out + b * 16);
  }
}

// Balanced variant: the (sample, tile) space is cut into gridDim.x CONTIGUOUS ranges of tiles (tile = blockDim * UNR items of
// one sample), one CTA per range and exactly as many CTAs as are resident at once -- no partial second wave (the 2-D grid above
// ran 2.7 waves at 3 CTAs / SM and lost 10-30 % to the tail at C >= 96).  A CTA crosses at most a few sample boundaries; there
// it flushes its statistics and reloads the per-channel scale / shift.
template <int UNR, bool RES2>
__global__ void __launch_bounds__(256) gn_apply_ranges_kernel(const GnArgs a, int64_t tiles_per_sample, int64_t tiles_total) {
  __shared__ float s_scale[256], s_shift[256];
  __shared__ float s_part[256][4];
  const int C = a.C, cg = C >> 3;
  const int c0 = (threadIdx.x * 8) % C;   // blockDim * 8 is a multiple of C and tiles start on multiples of blockDim
  const int64_t t_begin = tiles_total * blockIdx.x / gridDim.x, t_end = tiles_total * (blockIdx.x + 1) / gridDim.x;
  const int64_t items = a.per_sample / 8;
  float sc[8], sh[8];
  float sa1 = 0.f, sa2 = 0.f, sb1 = 0.f, sb2 = 0.f;
  int cur_b = -1;
  const uint4 *xp = nullptr, *rp = nullptr, *r2p = nullptr;
  uint4* yp = nullptr;

  __shared__ unsigned char s_grp[256][2];
  s_grp[threadIdx.x][0] = (unsigned char)(c0 / cg); s_grp[threadIdx.x][1] = (unsigned char)((c0 + 4) / cg);
  auto flush = [&](int b) {
    gn_block_stats_flush(s_part, s_grp, sa1, sa2, sb1, sb2, a.stats_out + b * 16);
    sa1 = sa2 = sb1 = sb2 = 0.f;
  };

  // (sample, tile inside the sample) advance incrementally: a 64-bit division per tile cost the loop ~10 % at four items per thread
  int b = (int)(t_begin / tiles_per_sample);
  int64_t tin = t_begin - (int64_t)b * tiles_per_sample;
  for (int64_t tile = t_begin; tile < t_end; ++tile, ++tin) {
    if (tin == tiles_per_sample) { tin = 0; ++b; }
    if (b != cur_b) {   // uniform across the CTA
      if (cur_b >= 0 && a.mode == 2 && a.stats_out) flush(cur_b);
      __syncthreads();
      if (threadIdx.x < C) {
        const int c = threadIdx.x, g = c / cg;
        const double n = (double)(a.per_sample / 8);
        const double mean = a.stats_in[(b * 8 + g) * 2] / n;
        double var = a.stats_in[(b * 8 + g) * 2 + 1] / n - mean * mean;
        if (var < 0) var = 0;
        const float rstd = (float)(1.0 / sqrt(var + (double)a.eps));
        const float scv = a.gamma[c] * rstd;
        s_scale[c] = scv;
        s_shift[c] = (a.beta ? a.beta[c] : 0.f) - (float)mean * scv;
      }
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 8; ++j) { sc[j] = s_scale[c0 + j]; sh[j] = s_shift[c0 + j]; }
      xp = reinterpret_cast<const uint4*>(a.x + (int64_t)b * a.per_sample);
      rp = a.res ? reinterpret_cast<const uint4*>(a.res + (int64_t)b * a.per_sample) : nullptr;
      r2p = a.res2 ? reinterpret_cast<const uint4*>(a.res2 + (int64_t)b * a.per_sample) : nullptr;
      yp = reinterpret_cast<uint4*>(a.y + (int64_t)b * a.per_sample);
      cur_b = b;
    }
    const int64_t base = tin * ((int64_t)blockDim.x * UNR) + threadIdx.x;
    uint4 xv[UNR], rv[UNR], qv[RES2 ? UNR : 1];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      xv[u] = ldg_nc_v4(xp + base + u * (int64_t)blockDim.x);
      if (a.mode == 2) rv[u] = ldg_nc_v4(rp + base + u * (int64_t)blockDim.x);
      if (RES2) qv[u] = ldg_nc_v4(r2p + base + u * (int64_t)blockDim.x);
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const uint32_t xw[4] = {xv[u].x, xv[u].y, xv[u].z, xv[u].w};
      float f[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 t = unpack_bf16x2(xw[j]);
        if (a.pre_silu) { t.x = silu_fast(t.x); t.y = silu_fast(t.y); }
        f[2 * j] = fmaf(t.x, sc[2 * j], sh[2 * j]);
        f[2 * j + 1] = fmaf(t.y, sc[2 * j + 1], sh[2 * j + 1]);
      }
      if (a.mode == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = silu_fast(f[j]);
      } else if (a.mode == 2) {
        const uint32_t rw[4] = {rv[u].x, rv[u].y, rv[u].z, rv[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 t = unpack_bf16x2(rw[j]);
          f[2 * j] += t.x; f[2 * j + 1] += t.y;
        }
        if (RES2) {
          const uint32_t qw[4] = {qv[u].x, qv[u].y, qv[u].z, qv[u].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 t = unpack_bf16x2(qw[j]);
            f[2 * j] += t.x; f[2 * j + 1] += t.y;
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) { sa1 += f[j]; sa2 = fmaf(f[j], f[j], sa2); sb1 += f[4 + j]; sb2 = fmaf(f[4 + j], f[4 + j], sb2); }
      }
      stg_v4(yp + base + u * (int64_t)blockDim.x,
             make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7])));
    }
  }
  (void)items;
  if (cur_b >= 0 && a.mode == 2 && a.stats_out) flush(cur_b);
}

template <int UNR, bool RES2>
static int gn_ranges_launch(const GnArgs& a, int B, int threads, int64_t items, int waves_pct, cudaStream_t stream) {
  const int64_t tile = (int64_t)threads * UNR;
  const int64_t tps = items / tile, total = tps * B;
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gn_apply_ranges_kernel<UNR, RES2>, threads, 0) != cudaSuccess || occ < 1) occ = 1;
  int64_t grid = (int64_t)ddimb_num_sms() * occ * waves_pct / 100;
  if (grid > total) grid = total;
  if (grid < 1) grid = 1;
  gn_apply_ranges_kernel<UNR, RES2><<<(unsigned)grid, threads, 0, stream>>>(a, tps, total);
  return 0;
}

extern "C" int ddimb_gn_apply(const void* x, const void* res, const void* res2, void* y, const double* stats_in, double* stats_out, const float* gamma,
                              const float* beta, int mode, int B, long long positions, int C, float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && y && stats_in && gamma && B >= 1 && positions >= 1);
  DDIMB_CHECK_ARG(C % 32 == 0 && C <= 256 && mode >= 0 && mode <= 4);
  // modes 3 / 4 (training): the stored tensor is a pre-activation; y = gn(silu(x)) / y = res + gn(silu(x))
  const int pre_silu = mode >= 3 ? 1 : 0;
  if (mode == 3) mode = 0;
  if (mode == 4) mode = 2;
  DDIMB_CHECK_ARG(mode != 2 || res);
  DDIMB_CHECK_ARG(res2 == nullptr || mode == 2);
  DDIMB_PROF(stream, "gn_apply", C, mode, 2.0 * B * positions * C * (mode == 2 ? (res2 ? 4 : 3) : 2), 0.0);
  GnArgs a;
  a.x = reinterpret_cast<const __nv_bfloat16*>(x); a.res = reinterpret_cast<const __nv_bfloat16*>(res);
  a.res2 = reinterpret_cast<const __nv_bfloat16*>(res2);
  a.y = reinterpret_cast<__nv_bfloat16*>(y);
  a.stats_in = stats_in; a.stats_out = stats_out; a.gamma = gamma; a.beta = beta;
  a.per_sample = positions * C; a.C = C; a.mode = mode; a.pre_silu = pre_silu; a.eps = eps;
  const int threads = (C % 96 == 0) ? 192 : 256;   // threads * 8 must be a multiple of C
  const int64_t items = a.per_sample / 8;
  int64_t bx = (items + threads * 4 - 1) / (threads * 4);
  constexpr int unr = 4;
  // measured (tools/ab_gn.sh, B = 64, bench shapes): mode 2 -- ranges kernel +13 % at C = 96, +3 % at 128, +31 % at 192, but the
  // 2-D grid stays 4 % ahead at C = 32 / 64 (tensors large enough to hide its tail); modes 0 / 1 -- ranges +12 % (C = 64) to
  // +31 % (C = 96, 192); an unroll of 8 is no better than 4 anywhere
  const bool ranges = (mode != 2 || C >= 96);
  if (ranges && items % ((int64_t)threads * unr) == 0 && items / ((int64_t)threads * unr) >= 1) {
    if (a.res2) gn_ranges_launch<unr, true>(a, B, threads, items, 100, stream);
    else gn_ranges_launch<unr, false>(a, B, threads, items, 100, stream);   // balanced contiguous ranges, one resident wave
  } else {
    int64_t cap = (int64_t)ddimb_num_sms() * 8 / B + 1;
    if (bx > cap) bx = cap;
    if (bx < 1) bx = 1;
    if (a.res2) gn_apply_kernel<unr, true><<<dim3((unsigned)bx, B), threads, 0, stream>>>(a);
    else gn_apply_kernel<unr, false><<<dim3((unsigned)bx, B), threads, 0, stream>>>(a);
  }
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ small fp32 Linear (temb MLP)
// reference models/diffusion.py:107-120 (BetaEmbedding): out[r][n] = act(sum_k in[idx ? idx[r] : r][k] * w[n][k] + bias[n]).
// One warp per output feature, kept in fp32 for parity.  The (activated) input rows are staged in shared memory once per CTA: with a
// SiLU on the input the first version evaluated it per (output feature, row, k) -- 72 M exact exponentials for the training step's
// [32 x 512] x [512 x 4416] layer, 171 us.
__global__ void __launch_bounds__(256) linear_small_kernel(const float* __restrict__ in, const int64_t* __restrict__ idx,
                                                          const float* __restrict__ w, const float* __restrict__ bias,
                                                          float* __restrict__ out, int rows, int K, int N, int act) {
  extern __shared__ float s_in[];                          // [<= 64 rows][K]: blockIdx.y = chunk of 64 rows
  const int r0 = blockIdx.y * 64;
  rows = min(64, rows - r0);
  out += (int64_t)r0 * N;
  for (int i = threadIdx.x; i < rows * K; i += blockDim.x) {
    const int r = i / K, k = i - r * K;
    float iv = in[(idx ? idx[r0 + r] : (int64_t)(r0 + r)) * K + k];
    if (act & 2) iv = silu_exact(iv);
    s_in[i] = iv;
  }
  __syncthreads();
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float wr[16];   // K <= 512
  const int per = K >> 5;
#pragma unroll
  for (int j = 0; j < 16; ++j) wr[j] = j < per ? w[(int64_t)n * K + j * 32 + lane] : 0.f;
  const float bv = bias[n];
  for (int r = 0; r < rows; ++r) {
    const float* ip = s_in + r * K;
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (j < per) acc = fmaf(wr[j], ip[j * 32 + lane], acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      float v = acc + bv;
      if (act & 1) v = silu_exact(v);
      out[(int64_t)r * N + n] = v;
    }
  }
}

extern "C" int ddimb_linear_small(const float* in, const long long* idx, const float* w, const float* bias, float* out, int rows, int K,
                                  int N, int act, cudaStream_t stream) {
  DDIMB_CHECK_ARG(in && w && bias && out && rows >= 1 && rows <= 64 * 65535 && K % 32 == 0 && K <= 512 && N >= 1);
  DDIMB_PROF(stream, "linear_small", N, K, 4.0 * ((double)N * K + (double)rows * (K + N)), 2.0 * rows * N * K);
  const size_t smem = (size_t)(rows < 64 ? rows : 64) * K * sizeof(float);
  DDIMB_SMEM_ATTR(linear_small_kernel, 64 * 512 * sizeof(float));
  linear_small_kernel<<<dim3((N + 7) / 8, (rows + 63) / 64), 256, smem, stream>>>(in, reinterpret_cast<const int64_t*>(idx), w, bias, out, rows, K, N, act);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// out[r][:] = table[idx[r]][:] (fp32, width a multiple of 4): the per-step lookup into the precomputed timestep-embedding table
// (models/diffusion.py:107-120 evaluated once for t = 0..999 in eval mode; during sampling every row of t is the same step).
__global__ void __launch_bounds__(256) gather_rows_kernel(const float4* __restrict__ table, const int64_t* __restrict__ idx, float4* __restrict__ out,
                                                          int width4, int table_rows) {
  const int r = blockIdx.y;
  int64_t src = idx[r];
  src = src < 0 ? 0 : (src >= table_rows ? table_rows - 1 : src);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < width4; i += gridDim.x * blockDim.x) out[(int64_t)r * width4 + i] = table[src * width4 + i];
}

extern "C" int ddimb_gather_rows(const float* table, const long long* idx, float* out, int rows, int width, int table_rows, cudaStream_t stream) {
  DDIMB_CHECK_ARG(table && idx && out && rows >= 1 && rows <= 65535 && width >= 4 && width % 4 == 0 && table_rows >= 1);
  DDIMB_PROF(stream, "gather_rows", width, rows, 8.0 * rows * width, 0.0);
  gather_rows_kernel<<<dim3((width / 4 + 255) / 256, rows), 256, 0, stream>>>(reinterpret_cast<const float4*>(table), reinterpret_cast<const int64_t*>(idx),
                                                                                reinterpret_cast<float4*>(out), width / 4, table_rows);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ token path
// reference models/diffusion.py:273-275 + :130-145: permute(0,2,1,3) / reshape to tokens, + positional table, LayerNorm.
// x is NHWC bf16 [B][S][W][C]; token feature index is c*W + f (the reference flattens [C, F]); te is fp32 [>=S][C*W].
// One CTA of C threads per token; out bf16 [B*S][C*W].
template <int W>
__global__ void __launch_bounds__(256) tokens_ln_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ te,
                                                        const float* __restrict__ gamma, const float* __restrict__ beta,
                                                        __nv_bfloat16* __restrict__ out, int S, int C, float eps) {
  const int row = blockIdx.x, s = row % S, c = threadIdx.x;
  const int D = C * W;
  __shared__ float red[2][8];
  float v[W];
  const __nv_bfloat16* xp = x + (int64_t)row * W * C;
  const float* tp = te + (int64_t)s * D + c * W;
  float sum = 0.f;
#pragma unroll
  for (int f = 0; f < W; ++f) { v[f] = __bfloat162float(xp[f * C + c]) + tp[f]; sum += v[f]; }
  sum = warp_sum(sum);
  if ((c & 31) == 0) red[0][c >> 5] = sum;
  __syncthreads();
  float tot = 0.f;
  for (int i = 0; i < (C >> 5); ++i) tot += red[0][i];
  const float mean = tot / D;
  float sq = 0.f;
#pragma unroll
  for (int f = 0; f < W; ++f) { const float d = v[f] - mean; sq = fmaf(d, d, sq); }
  sq = warp_sum(sq);
  if ((c & 31) == 0) red[1][c >> 5] = sq;
  __syncthreads();
  float tsq = 0.f;
  for (int i = 0; i < (C >> 5); ++i) tsq += red[1][i];
  const float rstd = rsqrtf(tsq / D + eps);
  uint32_t o[W / 2];
#pragma unroll
  for (int f = 0; f < W; f += 2) {
    const float a = (v[f] - mean) * rstd * gamma[c * W + f] + beta[c * W + f];
    const float b2 = (v[f + 1] - mean) * rstd * gamma[c * W + f + 1] + beta[c * W + f + 1];
    o[f >> 1] = pack_bf16x2(a, b2);
  }
  uint4* op = reinterpret_cast<uint4*>(out + (int64_t)row * D + c * W);
#pragma unroll
  for (int i = 0; i < W / 8; ++i) stg_v4(op + i, make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]));
}

extern "C" int ddimb_tokens_ln(const void* x, const float* te, const float* gamma, const float* beta, void* out, int B, int S, int W,
                               int C, float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && te && gamma && beta && out && W == 8 && C % 32 == 0 && C <= 256);
  DDIMB_PROF(stream, "tokens_ln", C, W, (double)B * S * W * C * 4.0 + (double)S * W * C * 4.0, 0.0);
  tokens_ln_kernel<8><<<B * S, C, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), te, gamma, beta,
                                                reinterpret_cast<__nv_bfloat16*>(out), S, C, eps);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// reference models/diffusion.py:277-279,284: tokens back to [B, C, S, W] and the first additive skip; also the
// GroupNorm statistics of the result.  tok is fp32 [B*S][C*W]; skip/out are NHWC bf16 [B][S][W][C].
template <int W>
__global__ void __launch_bounds__(256) tokens_to_nhwc_kernel(const float* __restrict__ tok, const __nv_bfloat16* __restrict__ skip,
                                                             __nv_bfloat16* __restrict__ out, double* __restrict__ stats, int S, int C) {
  const int row = blockIdx.x, b = row / S, c = threadIdx.x;
  const float* tp = tok + (int64_t)row * C * W + c * W;
  const __nv_bfloat16* sp = skip + (int64_t)row * W * C;
  __nv_bfloat16* op = out + (int64_t)row * W * C;
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int f = 0; f < W; ++f) {
    const float v = tp[f] + __bfloat162float(sp[f * C + c]);
    op[f * C + c] = __float2bfloat16_rn(v);
    s1 += v; s2 = fmaf(v, v, s2);
  }
  if (stats) {
    // C/8 channels per group; with C = 256 a warp is exactly one group, otherwise reduce in sub-warp segments
    const int cg = C >> 3;
    for (int o = 1; o < cg && o < 32; o <<= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
    if ((c % (cg < 32 ? cg : 32)) == 0) {
      const int g = c / cg;
      atomicAdd(stats + b * 16 + 2 * g, (double)s1);
      atomicAdd(stats + b * 16 + 2 * g + 1, (double)s2);
    }
  }
}

extern "C" int ddimb_tokens_to_nhwc(const float* tok, const void* skip, void* out, double* stats, int B, int S, int W, int C,
                                    cudaStream_t stream) {
  DDIMB_CHECK_ARG(tok && skip && out && W == 8 && C % 32 == 0 && C <= 256 && ((C >> 3) & ((C >> 3) - 1)) == 0);
  DDIMB_PROF(stream, "tokens_to_nhwc", C, W, (double)B * S * W * C * 8.0, 0.0);
  tokens_to_nhwc_kernel<8><<<B * S, C, 0, stream>>>(tok, reinterpret_cast<const __nv_bfloat16*>(skip), reinterpret_cast<__nv_bfloat16*>(out),
                                                    stats, S, C);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ LayerNorm (+ residual)
// reference hf modeling_fnet.py:180-188, 226-231: y = LN(a + b) over H = 512, eps 1e-6.  One warp per row.
// Writes the fp32 residual stream and its bf16 GEMM operand (hi plane, and the rounding residual lo plane when split).
__global__ void __launch_bounds__(256) layernorm512_kernel(const float* __restrict__ a, const float* __restrict__ b2,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           float* __restrict__ y, __nv_bfloat16* __restrict__ ybf, int64_t ldy,
                                                           int write_lo, int rows, float eps) {
  constexpr int H = 512;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float4* ap = reinterpret_cast<const float4*>(a + (int64_t)row * H);
  const float4* bp = b2 ? reinterpret_cast<const float4*>(b2 + (int64_t)row * H) : nullptr;
  float4 v[4];
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[j] = ap[j * 32 + lane];
    if (bp) { const float4 t = bp[j * 32 + lane]; v[j].x += t.x; v[j].y += t.y; v[j].z += t.z; v[j].w += t.w; }
    sum += v[j].x + v[j].y + v[j].z + v[j].w;
  }
  const float mean = warp_sum(sum) / H;
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float d0 = v[j].x - mean, d1 = v[j].y - mean, d2 = v[j].z - mean, d3 = v[j].w - mean;
    sq += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
  }
  const float rstd = rsqrtf(warp_sum(sq) / H + eps);
  const float4* gp = reinterpret_cast<const float4*>(gamma);
  const float4* bt = reinterpret_cast<const float4*>(beta);
  float4* yp = reinterpret_cast<float4*>(y + (int64_t)row * H);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float4 g = gp[j * 32 + lane], be = bt[j * 32 + lane];
    float4 o;
    o.x = (v[j].x - mean) * rstd * g.x + be.x; o.y = (v[j].y - mean) * rstd * g.y + be.y;
    o.z = (v[j].z - mean) * rstd * g.z + be.z; o.w = (v[j].w - mean) * rstd * g.w + be.w;
    yp[j * 32 + lane] = o;
    const uint32_t h0 = pack_bf16x2(o.x, o.y), h1 = pack_bf16x2(o.z, o.w);
    uint2* hp = reinterpret_cast<uint2*>(ybf + (int64_t)row * ldy + (j * 32 + lane) * 4);
    *hp = make_uint2(h0, h1);
    if (write_lo) {
      const float2 f0 = unpack_bf16x2(h0), f1 = unpack_bf16x2(h1);
      *reinterpret_cast<uint2*>(ybf + (int64_t)row * ldy + H + (j * 32 + lane) * 4) =
          make_uint2(pack_bf16x2(o.x - f0.x, o.y - f0.y), pack_bf16x2(o.z - f1.x, o.w - f1.y));
    }
  }
}

extern "C" int ddimb_layernorm512(const float* a, const float* b, const float* gamma, const float* beta, float* y, void* ybf,
                                  long long ldy, int write_lo, int rows, float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(a && gamma && beta && y && ybf && rows >= 1 && ldy % 4 == 0 && ldy >= (write_lo ? 1024 : 512));
  DDIMB_PROF(stream, "layernorm512", write_lo, b ? 1 : 0, 512.0 * rows * (4.0 * (b ? 2 : 1) + 4.0 + (write_lo ? 4.0 : 2.0)), 0.0);
  layernorm512_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(a, b, gamma, beta, y, reinterpret_cast<__nv_bfloat16*>(ybf), ldy, write_lo, rows, eps);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// fp32 [rows][H] -> bf16 hi (and lo) planes in a [rows][ldy] operand buffer (projection output feeding the first FNet layer)
__global__ void __launch_bounds__(256) split_bf16_kernel(const float4* __restrict__ a, __nv_bfloat16* __restrict__ ybf, int64_t ldy,
                                                         int H, int write_lo, int64_t n4) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 o = a[i];
    const int64_t row = (i * 4) / H, col = (i * 4) % H;
    const uint32_t h0 = pack_bf16x2(o.x, o.y), h1 = pack_bf16x2(o.z, o.w);
    *reinterpret_cast<uint2*>(ybf + row * ldy + col) = make_uint2(h0, h1);
    if (write_lo) {
      const float2 f0 = unpack_bf16x2(h0), f1 = unpack_bf16x2(h1);
      *reinterpret_cast<uint2*>(ybf + row * ldy + H + col) = make_uint2(pack_bf16x2(o.x - f0.x, o.y - f0.y), pack_bf16x2(o.z - f1.x, o.w - f1.y));
    }
  }
}

extern "C" int ddimb_split_bf16(const float* a, void* ybf, long long ldy, int H, int write_lo, long long rows, cudaStream_t stream) {
  DDIMB_CHECK_ARG(a && ybf && rows >= 1 && H % 4 == 0 && ldy % 4 == 0);
  DDIMB_PROF(stream, "split_bf16", H, write_lo, (double)rows * H * (4.0 + (write_lo ? 4.0 : 2.0)), 0.0);
  const int64_t n4 = rows * H / 4;
  int64_t blocks = (n4 + 255) / 256;
  const int64_t cap = (int64_t)ddimb_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  split_bf16_kernel<<<(int)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(a), reinterpret_cast<__nv_bfloat16*>(ybf), ldy, H, write_lo, n4);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}
