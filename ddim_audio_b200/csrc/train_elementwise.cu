// Backward-pass kernels of the HBM-bound layers (autograd of reference models/diffusion.py:12-56,81-145,189-208 and
// hf modeling_fnet.py:180-231): GroupNorm / SiLU, LayerNorm, gelu_new, token repacks, the temb MLP, conv_in / conv_out
// parameter gradients, dropout.  bf16 NHWC gradients, fp32 math, fp64 cross-block accumulation where a sum spans a sample.
#include "../../include/ddim_b200.h"
#include "common.cuh"

using namespace ddimb;

namespace {

__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_approx(0.5f * x), 0.5f); }
// d/dx silu(x) = s (1 + x (1 - s)), s = sigmoid(x)
__device__ __forceinline__ float dsilu_fast(float x) {
  const float s = sigmoid_fast(x);
  return s * fmaf(x, 1.0f - s, 1.0f);
}
__device__ __forceinline__ float dsilu_exact(float x) {
  const float s = 1.0f / (1.0f + __expf(-x));
  return s * fmaf(x, 1.0f - s, 1.0f);
}
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = unpack_bf16x2(w[j]);
    f[2 * j] = t.x; f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}

// Per-(sample, channel) sums of a thread's 8 channels: every thread of a block owns the 8 channels starting at
// (threadIdx.x * 8) % C for all items it visits (blockDim * 8 is a multiple of C).  Fixed-order block reduction, then one
// double atomicAdd per (channel, value) per block.
template <int NV>
__device__ __forceinline__ void block_channel_sums(float (&acc)[NV][8], double* dst /* [C][NV] of this sample */, int C, float* s_part) {
  // s_part: [blockDim][NV * 8]
  float* mine = s_part + threadIdx.x * (NV * 8);
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int j = 0; j < 8; ++j) mine[v * 8 + j] = acc[v][j];
  __syncthreads();
  const int chunks = C >> 3;   // distinct 8-channel chunks; thread t owns chunk t % chunks
  for (int o = threadIdx.x; o < C * NV; o += blockDim.x) {
    const int c = o / NV, v = o - c * NV;
    const int chunk = c >> 3, j = c & 7;
    float s = 0.f;
    for (int t = chunk; t < (int)blockDim.x; t += chunks) s += s_part[t * (NV * 8) + v * 8 + j];
    atomicAdd(dst + (int64_t)c * NV + v, (double)s);
  }
}

struct GnBwdArgs {
  const __nv_bfloat16* g;    // gradient at the GroupNorm output (type B) / at the SiLU output (type A)
  const __nv_bfloat16* z;    // type A: the GroupNorm input x;  type B: the pre-activation u (GroupNorm input is silu(u))
  const __nv_bfloat16* add;  // optional tensor added to the result (type A: the residual-path gradient)
  __nv_bfloat16* out;
  const double* stats;       // [B][8][2] (sum, sumsq) of the GroupNorm input
  double* R;                 // [B][C][2]: sum over positions of gh, gh * (GroupNorm input)
  double* sums_out;          // optional [B][C]: sum over positions of out (type B: d bias / d temb)
  const float* gamma;
  const float* beta;
  int64_t per_sample;
  int C;
  int type;                  // 0 = A: y = silu(gn(x));  1 = B: y = gn(silu(u))
  float eps;
};

__device__ __forceinline__ void gn_group_stats(const double* stats, int b, int g, double n, float eps, float& mean, float& rstd) {
  const double m = stats[(b * 8 + g) * 2] / n;
  double var = stats[(b * 8 + g) * 2 + 1] / n - m * m;
  if (var < 0) var = 0;
  mean = (float)m;
  rstd = (float)(1.0 / sqrt(var + (double)eps));
}

// Pass 1: R[b][c] = (sum gh, sum gh * zin) with  A: gh = g * silu'(gn(x)), zin = x;   B: gh = g, zin = silu(u)
__global__ void __launch_bounds__(256) gn_bwd_reduce_kernel(const GnBwdArgs a) {
  extern __shared__ float s_dyn[];
  __shared__ float s_scale[256], s_shift[256];
  const int b = blockIdx.y, C = a.C, cg = C >> 3;
  if (a.type == 0 && threadIdx.x < C) {
    const int c = threadIdx.x;
    float mean, rstd;
    gn_group_stats(a.stats, b, c / cg, (double)(a.per_sample / 8), a.eps, mean, rstd);
    const float sc = a.gamma[c] * rstd;
    s_scale[c] = sc;
    s_shift[c] = (a.beta ? a.beta[c] : 0.f) - mean * sc;
  }
  __syncthreads();
  const int c0 = (threadIdx.x * 8) % C;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = a.type == 0 ? s_scale[c0 + j] : 1.f; sh[j] = a.type == 0 ? s_shift[c0 + j] : 0.f; }
  const int64_t items = a.per_sample / 8;
  const uint4* gp = reinterpret_cast<const uint4*>(a.g + (int64_t)b * a.per_sample);
  const uint4* zp = reinterpret_cast<const uint4*>(a.z + (int64_t)b * a.per_sample);
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  // four items per iteration with every load issued before the first use (one item at a time left 32 bytes per thread in flight:
  // 0.65 of HBM at 3 resident CTAs per SM)
  const int64_t S = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < items; i0 += 4 * S) {
    uint4 gv[4], zv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * S;
      if (i < items) { gv[u] = ldg_nc_v4(gp + i); zv[u] = ldg_nc_v4(zp + i); }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (i0 + u * S >= items) break;
      float g[8], z[8];
      unpack8(gv[u], g);
      unpack8(zv[u], z);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float gh, zin;
        if (a.type == 0) { gh = g[j] * dsilu_fast(fmaf(z[j], sc[j], sh[j])); zin = z[j]; }
        else { gh = g[j]; zin = silu_fast(z[j]); }
        acc[0][j] += gh;
        acc[1][j] = fmaf(gh, zin, acc[1][j]);
      }
    }
  }
  block_channel_sums<2>(acc, a.R + (int64_t)b * C * 2, C, s_dyn);
}

// Pass 2: the GroupNorm input gradient  dzin = A_c gh + B_g zin + C_g  (A_c = rstd gamma_c, B_g = -rstd^2 k2, C_g = rstd (rstd mean k2 - k1),
// k1 = mean_group(gamma gh), k2 = mean_group(gamma gh zhat)), then  A: out = dzin + add;   B: out = dzin * silu'(u), sums of out.
__global__ void __launch_bounds__(256) gn_bwd_apply_kernel(const GnBwdArgs a) {
  extern __shared__ float s_dyn[];
  __shared__ float s_scale[256], s_shift[256], s_A[256], s_k[256][2], s_B[8], s_C[8];
  const int b = blockIdx.y, C = a.C, cg = C >> 3;
  const double n = (double)(a.per_sample / 8);
  if (threadIdx.x < C) {
    const int c = threadIdx.x;
    float mean, rstd;
    gn_group_stats(a.stats, b, c / cg, n, a.eps, mean, rstd);
    const float gm = a.gamma[c];
    const double r0 = a.R[((int64_t)b * C + c) * 2], r1 = a.R[((int64_t)b * C + c) * 2 + 1];
    s_k[c][0] = (float)(gm * r0 / n);
    s_k[c][1] = (float)(gm * (double)rstd * (r1 - (double)mean * r0) / n);
    s_A[c] = rstd * gm;
    s_scale[c] = gm * rstd;
    s_shift[c] = (a.beta ? a.beta[c] : 0.f) - mean * gm * rstd;
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    const int g = threadIdx.x;
    float k1 = 0.f, k2 = 0.f;
    for (int c = g * cg; c < (g + 1) * cg; ++c) { k1 += s_k[c][0]; k2 += s_k[c][1]; }
    float mean, rstd;
    gn_group_stats(a.stats, b, g, n, a.eps, mean, rstd);
    s_B[g] = -rstd * rstd * k2;
    s_C[g] = rstd * (rstd * mean * k2 - k1);
  }
  __syncthreads();
  const int c0 = (threadIdx.x * 8) % C;
  float sc[8], sh[8], cA[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = s_scale[c0 + j]; sh[j] = s_shift[c0 + j]; cA[j] = s_A[c0 + j]; }
  // group width is a multiple of 4: channels c0..c0+3 and c0+4..c0+7 each lie in one group
  const float cB0 = s_B[c0 / cg], cC0 = s_C[c0 / cg], cB1 = s_B[(c0 + 4) / cg], cC1 = s_C[(c0 + 4) / cg];
  const int64_t items = a.per_sample / 8;
  const uint4* gp = reinterpret_cast<const uint4*>(a.g + (int64_t)b * a.per_sample);
  const uint4* zp = reinterpret_cast<const uint4*>(a.z + (int64_t)b * a.per_sample);
  const uint4* ap = a.add ? reinterpret_cast<const uint4*>(a.add + (int64_t)b * a.per_sample) : nullptr;
  uint4* op = reinterpret_cast<uint4*>(a.out + (int64_t)b * a.per_sample);
  float acc[1][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[0][j] = 0.f;
  const int64_t S = (int64_t)gridDim.x * blockDim.x;     // four items per iteration, loads first (see gn_bwd_reduce_kernel)
  for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < items; i0 += 4 * S) {
    uint4 gv[4], zv[4], av[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * S;
      if (i < items) {
        gv[u] = ldg_nc_v4(gp + i);
        zv[u] = ldg_nc_v4(zp + i);
        if (ap) av[u] = ldg_nc_v4(ap + i);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t i = i0 + u * S;
      if (i >= items) break;
      float g[8], z[8], ad[8], o[8];
      unpack8(gv[u], g);
      unpack8(zv[u], z);
      if (ap) unpack8(av[u], ad);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float cB = j < 4 ? cB0 : cB1, cC = j < 4 ? cC0 : cC1;
        if (a.type == 0) {
          const float gh = g[j] * dsilu_fast(fmaf(z[j], sc[j], sh[j]));
          o[j] = fmaf(cA[j], gh, fmaf(cB, z[j], cC));
        } else {
          const float sg = sigmoid_fast(z[j]);              // one tanh serves silu(u) and silu'(u)
          const float zin = z[j] * sg;
          o[j] = fmaf(cA[j], g[j], fmaf(cB, zin, cC)) * (sg * fmaf(z[j], 1.0f - sg, 1.0f));
        }
        if (ap) o[j] += ad[j];
      }
      const uint4 ov = pack8(o);
      stg_v4(op + i, ov);
      if (a.sums_out) {
        float r[8];
        unpack8(ov, r);   // sums of the STORED (bf16) values: what the weight-gradient GEMM of the same tensor sees
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[0][j] += r[j];
      }
    }
  }
  if (a.sums_out) block_channel_sums<1>(acc, a.sums_out + (int64_t)b * C, C, s_dyn);
}

// d gamma[c] += sum_b rstd (R1 - mean R0);  d beta[c] += sum_b R0.  A warp per channel: lanes take samples b = lane, lane + 32, ...
// (each term costs an fp64 division and square root: the first version walked the samples serially, 20-30 us per call x 96 calls per
// train step), joined by a fixed shuffle tree.
__global__ void __launch_bounds__(256) gn_bwd_param_kernel(const double* __restrict__ R, const double* __restrict__ stats, float* __restrict__ dgamma,
                                                           float* __restrict__ dbeta, int B, int C, double n, float eps) {
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  const int g = c / (C >> 3);
  double dg = 0, db = 0;
  for (int b = lane; b < B; b += 32) {
    float mean, rstd;
    gn_group_stats(stats, b, g, n, eps, mean, rstd);
    const double r0 = R[((int64_t)b * C + c) * 2], r1 = R[((int64_t)b * C + c) * 2 + 1];
    dg += (double)rstd * (r1 - (double)mean * r0);
    db += r0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    dg += __shfl_xor_sync(0xffffffffu, dg, o);
    db += __shfl_xor_sync(0xffffffffu, db, o);
  }
  if (lane == 0) {
    dgamma[c] += (float)dg;
    if (dbeta) dbeta[c] += (float)db;
  }
}

// CTAs per sample: at most ONE resident wave over all samples (occ = CTAs of this kernel an SM holds: 3 at 78 registers).  The first
// version sized the grid for 8 CTAs per SM: 2.7 waves at x[32,2,1024,256], the last one a third full.
template <typename K>
int gn_bwd_grid(K kernel, int64_t per_sample, int B, int threads, size_t smem) {
  const int64_t items = per_sample / 8;
  int64_t bx = (items + threads * 4 - 1) / (threads * 4);
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, smem) != cudaSuccess || occ < 1) occ = 1;
  int64_t cap = (int64_t)ddimb_num_sms() * occ / B;
  if (cap < 1) cap = 1;
  if (bx > cap) bx = cap;
  return bx < 1 ? 1 : (int)bx;
}

}  // namespace

// type 0 (A): y = silu(gn(x)): g = dL/dy, z = x.   type 1 (B): y = gn(silu(u)): g = dL/dy, z = u.
// R is double [B][C][2], zeroed by the caller.
extern "C" int ddimb_gn_bwd_reduce(const void* g, const void* z, const double* stats, const float* gamma, const float* beta, double* R,
                                   int type, int B, long long positions, int C, float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(g && z && stats && gamma && R && B >= 1 && positions >= 1 && C % 32 == 0 && C <= 256 && (type == 0 || type == 1));
  DDIMB_PROF(stream, "gn_bwd_reduce", C, type, 4.0 * B * positions * C, 0.0);
  GnBwdArgs a = {};
  a.g = reinterpret_cast<const __nv_bfloat16*>(g); a.z = reinterpret_cast<const __nv_bfloat16*>(z);
  a.stats = stats; a.R = R; a.gamma = gamma; a.beta = beta; a.per_sample = positions * C; a.C = C; a.type = type; a.eps = eps;
  const int threads = (C % 96 == 0) ? 192 : 256;
  gn_bwd_reduce_kernel<<<dim3(gn_bwd_grid(gn_bwd_reduce_kernel, a.per_sample, B, threads, threads * 16 * sizeof(float)), B), threads, threads * 16 * sizeof(float), stream>>>(a);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// out = gradient w.r.t. x (type A, + add) or w.r.t. u (type B); sums_out (nullable, double [B][C], zeroed by the caller)
// receives the per-(sample, channel) sums of out.
extern "C" int ddimb_gn_bwd_apply(const void* g, const void* z, const void* add, void* out, const double* stats, const double* R,
                                  const float* gamma, const float* beta, double* sums_out, int type, int B, long long positions, int C,
                                  float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(g && z && out && stats && gamma && R && B >= 1 && positions >= 1 && C % 32 == 0 && C <= 256 && (type == 0 || type == 1));
  DDIMB_PROF(stream, "gn_bwd_apply", C, type, 2.0 * B * positions * C * (add ? 4 : 3), 0.0);
  GnBwdArgs a = {};
  a.g = reinterpret_cast<const __nv_bfloat16*>(g); a.z = reinterpret_cast<const __nv_bfloat16*>(z);
  a.add = reinterpret_cast<const __nv_bfloat16*>(add); a.out = reinterpret_cast<__nv_bfloat16*>(out);
  a.stats = stats; a.R = const_cast<double*>(R); a.sums_out = sums_out; a.gamma = gamma; a.beta = beta;
  a.per_sample = positions * C; a.C = C; a.type = type; a.eps = eps;
  const int threads = (C % 96 == 0) ? 192 : 256;
  gn_bwd_apply_kernel<<<dim3(gn_bwd_grid(gn_bwd_apply_kernel, a.per_sample, B, threads, threads * 8 * sizeof(float)), B), threads, threads * 8 * sizeof(float), stream>>>(a);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// dgamma / dbeta (fp32 [C], ACCUMULATED) of one GroupNorm from R and the forward statistics; dbeta nullable (norm[2] has no bias).
extern "C" int ddimb_gn_bwd_param(const double* R, const double* stats, float* dgamma, float* dbeta, int B, long long positions, int C,
                                  float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(R && stats && dgamma && B >= 1 && C % 8 == 0);
  DDIMB_PROF(stream, "gn_bwd_param", C, 0, 16.0 * B * C, 0.0);
  gn_bwd_param_kernel<<<(C + 7) / 8, 256, 0, stream>>>(R, stats, dgamma, dbeta, B, C, (double)positions * (C / 8), eps);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ column sums (bias gradients)
// out[c] += sum_r x[r][c]; x is bf16 or fp32 [rows][ld].  Grid-stride over row blocks, one fp32 atomicAdd per column per block.
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, int64_t ld, int64_t rows, int C, float* __restrict__ out,
                                                     int rows_per_block) {
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  for (int c = blockIdx.y * blockDim.x + threadIdx.x; c < C; c += gridDim.y * blockDim.x) {
    float s = 0.f;
    for (int64_t r = r0; r < r1; ++r) {
      if constexpr (sizeof(T) == 2) s += __bfloat162float(x[r * ld + c]);
      else s += x[r * ld + c];
    }
    atomicAdd(out + c, s);
  }
}

// Contiguous bf16 [rows][C] with C in {32, 64, 96, 128, 192, 256}: the tensor is walked as a flat vector of 16-byte chunks; the block
// size (256, or 192 when 96 divides C) makes a thread's stride a multiple of C, so a thread always meets the same 8 columns and keeps
// 8 register sums; shared-memory atomics fold the block, one global atomicAdd per column per block.  (r02: the generic kernel above
// used 32 of 256 lanes at C = 32 with 2-byte loads: 0.04 of HBM on the bias gradients of the level-0 Upsample.)
template <int THREADS>
__global__ void __launch_bounds__(THREADS) colsum_flat_bf16_kernel(const uint4* __restrict__ x, int64_t nchunks, int C, float* __restrict__ out) {
  __shared__ float s_acc[256];
  for (int c = threadIdx.x; c < C; c += THREADS) s_acc[c] = 0.f;
  __syncthreads();
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x; i < nchunks; i += (int64_t)gridDim.x * THREADS) {
    const uint4 v = ddimb::ldg_nc_v4(x + i);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = ddimb::unpack_bf16x2(w[j]);
      acc[2 * j] += f.x;
      acc[2 * j + 1] += f.y;
    }
  }
  const int c0 = (threadIdx.x * 8) % C;
#pragma unroll
  for (int j = 0; j < 8; ++j) atomicAdd(&s_acc[c0 + j], acc[j]);
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += THREADS) atomicAdd(out + c, s_acc[c]);
}

extern "C" int ddimb_colsum(const void* x, int is_bf16, long long ld, long long rows, int C, float* out, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && out && rows >= 1 && C >= 1 && ld >= C);
  DDIMB_PROF(stream, "colsum", C, is_bf16, (double)rows * C * (is_bf16 ? 2 : 4), 0.0);
  if (is_bf16 && ld == C && C % 32 == 0 && C <= 256 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const int64_t nchunks = rows * C / 8;
    const int threads = (C % 96 == 0) ? 192 : 256;               // threads * 8 is a multiple of C
    int64_t blocks = (nchunks + threads * 8 - 1) / (threads * 8);
    if (blocks > 8 * ddimb_num_sms()) blocks = 8 * ddimb_num_sms();
    if (threads == 192) colsum_flat_bf16_kernel<192><<<(unsigned)blocks, 192, 0, stream>>>(reinterpret_cast<const uint4*>(x), nchunks, C, out);
    else colsum_flat_bf16_kernel<256><<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(x), nchunks, C, out);
    DDIMB_LAUNCH_CHECK();
    return DDIMB_OK;
  }
  int rpb = (int)((rows + 4 * ddimb_num_sms() - 1) / (4 * ddimb_num_sms()));
  if (rpb < 8) rpb = 8;
  const dim3 grid((unsigned)((rows + rpb - 1) / rpb), (unsigned)((C + 255) / 256));
  if (is_bf16) colsum_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), ld, rows, C, out, rpb);
  else colsum_kernel<float><<<grid, 256, 0, stream>>>(reinterpret_cast<const float*>(x), ld, rows, C, out, rpb);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// double [B][C] per-sample sums -> fp32: out_b[b * ldb + c] += s (the temb slice gradient) and / or out_c[c] += sum_b s (a bias gradient)
__global__ void sums_scatter_kernel(const double* __restrict__ s, float* __restrict__ out_b, int64_t ldb, float* __restrict__ out_c, int B, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double tot = 0;
  for (int b = 0; b < B; ++b) {
    const double v = s[(int64_t)b * C + c];
    tot += v;
    if (out_b) out_b[(int64_t)b * ldb + c] += (float)v;
  }
  if (out_c) out_c[c] += (float)tot;
}

extern "C" int ddimb_sums_scatter(const double* sums, float* out_b, long long ldb, float* out_c, int B, int C, cudaStream_t stream) {
  DDIMB_CHECK_ARG(sums && (out_b || out_c) && B >= 1 && C >= 1);
  DDIMB_PROF(stream, "sums_scatter", C, 0, 8.0 * B * C, 0.0);
  sums_scatter_kernel<<<(C + 63) / 64, 64, 0, stream>>>(sums, out_b, ldb, out_c, B, C);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ LayerNorm(512) backward
// forward (hf modeling_fnet.py:180-188,226-231): y = LN(a + b).  dy = dy1 + dy2 (dy2 nullable).  One warp per row recomputes
// mean / rstd, writes ds = dL/d(a+b) as fp32 and as bf16 operand planes (hi | lo), accumulates dgamma / dbeta.
__global__ void __launch_bounds__(256) layernorm512_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b2,
                                                               const float* __restrict__ dy1, const float* __restrict__ dy2,
                                                               const float* __restrict__ gamma, float* __restrict__ ds,
                                                               __nv_bfloat16* __restrict__ dsbf, int64_t ldy, int write_lo,
                                                               float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, float eps) {
  constexpr int H = 512;
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const float4* gp = reinterpret_cast<const float4*>(gamma);
  float4 gm[4], dgacc[4], dbacc[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { gm[j] = gp[j * 32 + lane]; dgacc[j] = make_float4(0, 0, 0, 0); dbacc[j] = make_float4(0, 0, 0, 0); }
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const float4* ap = reinterpret_cast<const float4*>(a + (int64_t)row * H);
    const float4* bp = b2 ? reinterpret_cast<const float4*>(b2 + (int64_t)row * H) : nullptr;
    const float4* d1 = reinterpret_cast<const float4*>(dy1 + (int64_t)row * H);
    const float4* d2 = dy2 ? reinterpret_cast<const float4*>(dy2 + (int64_t)row * H) : nullptr;
    float4 v[4], d[4];
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[j] = ap[j * 32 + lane];
      if (bp) { const float4 t = bp[j * 32 + lane]; v[j].x += t.x; v[j].y += t.y; v[j].z += t.z; v[j].w += t.w; }
      d[j] = d1[j * 32 + lane];
      if (d2) { const float4 t = d2[j * 32 + lane]; d[j].x += t.x; d[j].y += t.y; d[j].z += t.z; d[j].w += t.w; }
      sum += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    const float mean = warp_sum(sum) / H;
    float sq = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[j].x -= mean; v[j].y -= mean; v[j].z -= mean; v[j].w -= mean;
      sq += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
    }
    const float rstd = rsqrtf(warp_sum(sq) / H + eps);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[j].x *= rstd; v[j].y *= rstd; v[j].z *= rstd; v[j].w *= rstd;   // v = xhat
      dgacc[j].x = fmaf(d[j].x, v[j].x, dgacc[j].x); dgacc[j].y = fmaf(d[j].y, v[j].y, dgacc[j].y);
      dgacc[j].z = fmaf(d[j].z, v[j].z, dgacc[j].z); dgacc[j].w = fmaf(d[j].w, v[j].w, dgacc[j].w);
      dbacc[j].x += d[j].x; dbacc[j].y += d[j].y; dbacc[j].z += d[j].z; dbacc[j].w += d[j].w;
      d[j].x *= gm[j].x; d[j].y *= gm[j].y; d[j].z *= gm[j].z; d[j].w *= gm[j].w;   // d = dxhat
      m1 += d[j].x + d[j].y + d[j].z + d[j].w;
      m2 += d[j].x * v[j].x + d[j].y * v[j].y + d[j].z * v[j].z + d[j].w * v[j].w;
    }
    m1 = warp_sum(m1) / H;
    m2 = warp_sum(m2) / H;
    float4* op = reinterpret_cast<float4*>(ds + (int64_t)row * H);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float4 o;
      o.x = rstd * (d[j].x - m1 - v[j].x * m2); o.y = rstd * (d[j].y - m1 - v[j].y * m2);
      o.z = rstd * (d[j].z - m1 - v[j].z * m2); o.w = rstd * (d[j].w - m1 - v[j].w * m2);
      op[j * 32 + lane] = o;
      if (dsbf) {
        const uint32_t h0 = pack_bf16x2(o.x, o.y), h1 = pack_bf16x2(o.z, o.w);
        *reinterpret_cast<uint2*>(dsbf + (int64_t)row * ldy + (j * 32 + lane) * 4) = make_uint2(h0, h1);
        if (write_lo) {
          const float2 f0 = unpack_bf16x2(h0), f1 = unpack_bf16x2(h1);
          *reinterpret_cast<uint2*>(dsbf + (int64_t)row * ldy + H + (j * 32 + lane) * 4) =
              make_uint2(pack_bf16x2(o.x - f0.x, o.y - f0.y), pack_bf16x2(o.z - f1.x, o.w - f1.y));
        }
      }
    }
  }
  // the block's warps are joined in shared memory first (fixed order), then one atomic per column and block: with an atomic per
  // column and WARP, 1024 rows meant 1024-way contention on each of the 1024 addresses (60 us for a 4 MB problem)
  __shared__ float s_red[8][2][H];
  const int w = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    reinterpret_cast<float4*>(&s_red[w][0][0])[j * 32 + lane] = dgacc[j];
    reinterpret_cast<float4*>(&s_red[w][1][0])[j * 32 + lane] = dbacc[j];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * H; i += blockDim.x) {
    const int which = i / H, c = i - which * H;
    float acc = 0.f;
    for (int k = 0; k < wpb; ++k) acc += s_red[k][which][c];
    atomicAdd((which ? dbeta : dgamma) + c, acc);
  }
}

extern "C" int ddimb_layernorm512_bwd(const float* a, const float* b, const float* dy1, const float* dy2, const float* gamma, float* ds,
                                      void* dsbf, long long ldy, int write_lo, float* dgamma, float* dbeta, int rows, float eps,
                                      cudaStream_t stream) {
  DDIMB_CHECK_ARG(a && dy1 && gamma && ds && dgamma && dbeta && rows >= 1);
  DDIMB_CHECK_ARG(!dsbf || (ldy % 4 == 0 && ldy >= (write_lo ? 1024 : 512)));
  DDIMB_PROF(stream, "layernorm512_bwd", write_lo, 0, 512.0 * rows * 4.0 * (2 + (b ? 1 : 0) + (dy2 ? 1 : 0)), 0.0);
  int blocks = (rows + 15) / 16;                 // two rows per warp at least: halves the atomics against one more 1 us row per warp
  if (blocks > 2 * ddimb_num_sms()) blocks = 2 * ddimb_num_sms();
  layernorm512_bwd_kernel<<<blocks, 256, 0, stream>>>(a, b, dy1, dy2, gamma, ds, reinterpret_cast<__nv_bfloat16*>(dsbf), ldy, write_lo, dgamma,
                                                      dbeta, rows, eps);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ gelu_new forward / backward (bf16)
// training keeps the pre-activation u (hf activations.py:59-66): act = gelu_new(u);  du = dact * gelu_new'(u)
__global__ void __launch_bounds__(256) gelu_kernel(const uint4* __restrict__ u, const uint4* __restrict__ dact, uint4* __restrict__ out, int64_t n8) {
  const float k = 0.7978845608028654f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
    float x[8], o[8];
    unpack8(u[i], x);
    if (dact == nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = gelu_new_fast(x[j]);
    } else {
      float g[8];
      unpack8(dact[i], g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float x2 = x[j] * x[j];
        const float th = tanh_approx(k * fmaf(0.044715f * x[j], x2, x[j]));
        const float d = 0.5f * (1.0f + th) + 0.5f * x[j] * (1.0f - th * th) * k * fmaf(3.0f * 0.044715f, x2, 1.0f);
        o[j] = g[j] * d;
      }
    }
    out[i] = pack8(o);
  }
}

extern "C" int ddimb_gelu(const void* u, const void* dact, void* out, long long n, cudaStream_t stream) {
  DDIMB_CHECK_ARG(u && out && n > 0 && n % 8 == 0);
  DDIMB_PROF(stream, "gelu", dact ? 1 : 0, 0, 2.0 * n * (dact ? 3 : 2), 0.0);
  int64_t blocks = (n / 8 + 255) / 256;
  if (blocks > 8LL * ddimb_num_sms()) blocks = 8LL * ddimb_num_sms();
  gelu_kernel<<<(int)blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(u), reinterpret_cast<const uint4*>(dact), reinterpret_cast<uint4*>(out), n / 8);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// fp32 (a [+ b]) -> bf16 hi (+ lo) operand planes [rows][ldy]; optional in-place-safe fp32 sum output
__global__ void __launch_bounds__(256) add_split_kernel(const float4* __restrict__ a, const float4* __restrict__ b, float4* __restrict__ sum_out,
                                                        __nv_bfloat16* __restrict__ ybf, int64_t ldy, int H, int write_lo, int64_t n4) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 o = a[i];
    if (b) { const float4 t = b[i]; o.x += t.x; o.y += t.y; o.z += t.z; o.w += t.w; }
    if (sum_out) sum_out[i] = o;
    const int64_t row = (i * 4) / H, col = (i * 4) % H;
    const uint32_t h0 = pack_bf16x2(o.x, o.y), h1 = pack_bf16x2(o.z, o.w);
    *reinterpret_cast<uint2*>(ybf + row * ldy + col) = make_uint2(h0, h1);
    if (write_lo) {
      const float2 f0 = unpack_bf16x2(h0), f1 = unpack_bf16x2(h1);
      *reinterpret_cast<uint2*>(ybf + row * ldy + H + col) = make_uint2(pack_bf16x2(o.x - f0.x, o.y - f0.y), pack_bf16x2(o.z - f1.x, o.w - f1.y));
    }
  }
}

extern "C" int ddimb_add_split_bf16(const float* a, const float* b, float* sum_out, void* ybf, long long ldy, int H, int write_lo, long long rows,
                                    cudaStream_t stream) {
  DDIMB_CHECK_ARG(a && ybf && rows >= 1 && H % 4 == 0 && ldy % 4 == 0);
  DDIMB_PROF(stream, "add_split_bf16", H, write_lo, (double)rows * H * (4.0 * (b ? 2 : 1) + (write_lo ? 4.0 : 2.0)), 0.0);
  const int64_t n4 = rows * H / 4;
  int64_t blocks = (n4 + 255) / 256;
  if (blocks > 8LL * ddimb_num_sms()) blocks = 8LL * ddimb_num_sms();
  add_split_kernel<<<(int)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(a), reinterpret_cast<const float4*>(b),
                                                    reinterpret_cast<float4*>(sum_out), reinterpret_cast<__nv_bfloat16*>(ybf), ldy, H, write_lo, n4);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ dropout (13 sites of the FNet branch)
// reference models/diffusion.py:129,144 and hf modeling_fnet.py:226-231 (nn.Dropout(0.1), training only).  Counter-based mask:
// keep(i) = hash(seed, i) >= p * 2^32; y = x * keep / (1 - p).  Backward applies the same call to the gradient.
__device__ __forceinline__ uint32_t mix32(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return (uint32_t)x;
}
__global__ void __launch_bounds__(256) dropout_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n, uint32_t thresh, float scale,
                                                      uint64_t seed, const unsigned long long* __restrict__ seed_dev) {
  if (seed_dev) seed += *seed_dev;     // a captured launch (CUDA graph) takes the step's seed from device memory
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = mix32(seed + (uint64_t)i * 0x9E3779B97F4A7C15ULL) >= thresh ? x[i] * scale : 0.f;
}

extern "C" int ddimb_dropout(const float* x, float* y, long long n, float p, unsigned long long seed, const unsigned long long* seed_dev,
                             cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && y && n > 0 && p >= 0.f && p < 1.f);
  DDIMB_PROF(stream, "dropout", 0, 0, 8.0 * n, 0.0);
  int64_t blocks = (n + 255) / 256;
  if (blocks > 8LL * ddimb_num_sms()) blocks = 8LL * ddimb_num_sms();
  dropout_kernel<<<(int)blocks, 256, 0, stream>>>(x, y, n, (uint32_t)((double)p * 4294967296.0), 1.0f / (1.0f - p), seed, seed_dev);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ token path backward
// forward (models/diffusion.py:273-275,130-142): v = permute(x)[row][c*W+f] + te[s]; tok = LN_D(v) (D = C*W).
// dtok fp32 [B*S][D] -> dx bf16 NHWC [B][S][W][C] (+ add: the gradient of the additive skip that reads the same tensor,
// models/diffusion.py:284), dgamma / dbeta accumulated.  One CTA of C threads per token, as the forward kernel.
template <int W>
__global__ void __launch_bounds__(256) tokens_ln_bwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ te,
                                                            const float* __restrict__ gamma, const float* __restrict__ dtok,
                                                            const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ dx,
                                                            float* __restrict__ dgamma, float* __restrict__ dbeta, int S, int C, float eps) {
  const int row = blockIdx.x, s = row % S, c = threadIdx.x;
  const int D = C * W;
  __shared__ float red[4][8];
  float v[W], d[W];
  const __nv_bfloat16* xp = x + (int64_t)row * W * C;
  const float* tp = te + (int64_t)s * D + c * W;
  const float* dp = dtok + (int64_t)row * D + c * W;
  float sum = 0.f;
#pragma unroll
  for (int f = 0; f < W; ++f) { v[f] = __bfloat162float(xp[f * C + c]) + tp[f]; sum += v[f]; d[f] = dp[f]; }
  const int nw = C >> 5;
  auto block_sum = [&](float val, int slot) {
    val = warp_sum(val);
    if ((c & 31) == 0) red[slot][c >> 5] = val;
    __syncthreads();
    float t = 0.f;
    for (int i = 0; i < nw; ++i) t += red[slot][i];
    return t;
  };
  const float mean = block_sum(sum, 0) / D;
  float sq = 0.f;
#pragma unroll
  for (int f = 0; f < W; ++f) { v[f] -= mean; sq = fmaf(v[f], v[f], sq); }
  const float rstd = rsqrtf(block_sum(sq, 1) / D + eps);
  float m1 = 0.f, m2 = 0.f;
#pragma unroll
  for (int f = 0; f < W; ++f) {
    v[f] *= rstd;                                   // xhat
    atomicAdd(dgamma + c * W + f, d[f] * v[f]);
    atomicAdd(dbeta + c * W + f, d[f]);
    d[f] *= gamma[c * W + f];                       // dxhat
    m1 += d[f]; m2 = fmaf(d[f], v[f], m2);
  }
  m1 = block_sum(m1, 2) / D;
  m2 = block_sum(m2, 3) / D;
  __nv_bfloat16* op = dx + (int64_t)row * W * C;
  const __nv_bfloat16* ap = add ? add + (int64_t)row * W * C : nullptr;
#pragma unroll
  for (int f = 0; f < W; ++f) {
    float o = rstd * (d[f] - m1 - v[f] * m2);
    if (ap) o += __bfloat162float(ap[f * C + c]);
    op[f * C + c] = __float2bfloat16_rn(o);
  }
}

extern "C" int ddimb_tokens_ln_bwd(const void* x, const float* te, const float* gamma, const float* dtok, const void* add, void* dx,
                                   float* dgamma, float* dbeta, int B, int S, int W, int C, float eps, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && te && gamma && dtok && dx && dgamma && dbeta && W == 8 && C % 32 == 0 && C <= 256);
  DDIMB_PROF(stream, "tokens_ln_bwd", C, W, (double)B * S * W * C * (2.0 + 4.0 + 2.0 + (add ? 2.0 : 0.0)), 0.0);
  tokens_ln_bwd_kernel<8><<<B * S, C, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), te, gamma, dtok,
                                                    reinterpret_cast<const __nv_bfloat16*>(add), reinterpret_cast<__nv_bfloat16*>(dx), dgamma, dbeta, S, C, eps);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// backward of models/diffusion.py:277-279: gradient NHWC bf16 [B][S][W][C] -> token-major bf16 [B*S][C*W] (GEMM operand)
template <int W>
__global__ void __launch_bounds__(256) nhwc_to_tokens_kernel(const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ tok, int C) {
  const int row = blockIdx.x, c = threadIdx.x;
  const __nv_bfloat16* gp = g + (int64_t)row * W * C;
  float v[W];
#pragma unroll
  for (int f = 0; f < W; ++f) v[f] = __bfloat162float(gp[f * C + c]);
  uint32_t o[W / 2];
#pragma unroll
  for (int f = 0; f < W; f += 2) o[f >> 1] = pack_bf16x2(v[f], v[f + 1]);
  uint4* op = reinterpret_cast<uint4*>(tok + (int64_t)row * C * W + c * W);
#pragma unroll
  for (int i = 0; i < W / 8; ++i) stg_v4(op + i, make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]));
}

extern "C" int ddimb_nhwc_to_tokens(const void* g, void* tok, int B, int S, int W, int C, cudaStream_t stream) {
  DDIMB_CHECK_ARG(g && tok && W == 8 && C % 32 == 0 && C <= 256);
  DDIMB_PROF(stream, "nhwc_to_tokens", C, W, (double)B * S * W * C * 4.0, 0.0);
  nhwc_to_tokens_kernel<8><<<B * S, C, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(g), reinterpret_cast<__nv_bfloat16*>(tok), C);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ temb MLP backward (fp32, B rows)
// forward (models/diffusion.py:107-120) as run in training: z1 = L0(te[t]); z2 = L1(silu(z1)); temb = L2(silu(z2)).
// dw[n][k] += sum_r g[r][n] * f(in[row(r)][k]),  db[n] += sum_r g[r][n];  f = SiLU when in_act.
__global__ void __launch_bounds__(512) linear_small_wgrad_kernel(const float* __restrict__ g, const float* __restrict__ in,
                                                                const int64_t* __restrict__ idx, float* __restrict__ dw,
                                                                float* __restrict__ db, int rows, int K, int N, int in_act) {
  const int n = blockIdx.x;
  float bsum = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float acc = 0.f;
    for (int r = 0; r < rows; ++r) {
      float iv = in[(idx ? idx[r] : (int64_t)r) * K + k];
      if (in_act) iv = silu_exact(iv);
      acc = fmaf(g[(int64_t)r * N + n], iv, acc);
    }
    dw[(int64_t)n * K + k] += acc;
  }
  if (threadIdx.x == 0) {
    for (int r = 0; r < rows; ++r) bsum += g[(int64_t)r * N + n];
    db[n] += bsum;
  }
}

extern "C" int ddimb_linear_small_wgrad(const float* g, const float* in, const long long* idx, float* dw, float* db, int rows, int K, int N,
                                        int in_act, cudaStream_t stream) {
  DDIMB_CHECK_ARG(g && in && dw && db && rows >= 1 && K >= 1 && N >= 1);
  DDIMB_PROF(stream, "linear_small_wgrad", N, K, 4.0 * ((double)N * K + (double)rows * (K + N)), 2.0 * rows * N * K);
  linear_small_wgrad_kernel<<<N, K < 512 ? K : 512, 0, stream>>>(g, in, reinterpret_cast<const int64_t*>(idx), dw, db, rows, K, N, in_act);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// out[r][k] = (z ? silu'(z[r][k]) : 1) * sum_n g[r][n] * w[n][k].  A CTA owns 32 columns k (lane = k) and up to 8 rows; its warps
// split n, every w element is read ONCE per CTA and meets all the CTA's rows from registers (g is staged in shared memory, broadcast
// reads).  (The first version ran one CTA per row with a serial loop over n of dependent loads: 313 us at [32 x 4416] x [4416 x 512].)
constexpr int kLsdRows = 8, kLsdWarps = 16;
__global__ void __launch_bounds__(32 * kLsdWarps) linear_small_dgrad_kernel(const float* __restrict__ g, const float* __restrict__ w,
                                                                          const float* __restrict__ z, float* __restrict__ out, int rows, int K, int N) {
  extern __shared__ float s_g[];                           // [N][kLsdRows] (row-minor: one 32-byte broadcast read per n), then the warps' partials
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * 32 + lane, r0 = blockIdx.y * kLsdRows;
  const int nr = min(kLsdRows, rows - r0);
  for (int i = threadIdx.x; i < N * kLsdRows; i += blockDim.x) {
    const int n = i / kLsdRows, r = i - n * kLsdRows;
    s_g[i] = r < nr ? g[(int64_t)(r0 + r) * N + n] : 0.f;
  }
  __syncthreads();
  float acc[kLsdRows];
#pragma unroll
  for (int r = 0; r < kLsdRows; ++r) acc[r] = 0.f;
  if (k < K) {
    for (int n = warp; n < N; n += kLsdWarps) {
      const float wv = __ldg(w + (int64_t)n * K + k);
      const float4* gp = reinterpret_cast<const float4*>(s_g + n * kLsdRows);
#pragma unroll
      for (int q = 0; q < kLsdRows / 4; ++q) {
        const float4 gv = gp[q];
        acc[4 * q] = fmaf(gv.x, wv, acc[4 * q]); acc[4 * q + 1] = fmaf(gv.y, wv, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(gv.z, wv, acc[4 * q + 2]); acc[4 * q + 3] = fmaf(gv.w, wv, acc[4 * q + 3]);
      }
    }
  }
  __syncthreads();                                         // s_g is reused for the partial sums [warp][row][lane]
#pragma unroll
  for (int r = 0; r < kLsdRows; ++r) s_g[(warp * kLsdRows + r) * 32 + lane] = acc[r];
  __syncthreads();
  for (int i = threadIdx.x; i < kLsdRows * 32; i += blockDim.x) {
    const int r = i >> 5, l = i & 31, kk = blockIdx.x * 32 + l;
    if (r >= nr || kk >= K) continue;
    float sum = 0.f;
#pragma unroll
    for (int wq = 0; wq < kLsdWarps; ++wq) sum += s_g[(wq * kLsdRows + r) * 32 + l];   // fixed order
    if (z) sum *= dsilu_exact(z[(int64_t)(r0 + r) * K + kk]);
    out[(int64_t)(r0 + r) * K + kk] = sum;
  }
}

extern "C" int ddimb_linear_small_dgrad(const float* g, const float* w, const float* z, float* out, int rows, int K, int N, cudaStream_t stream) {
  DDIMB_CHECK_ARG(g && w && out && rows >= 1 && K >= 1 && N >= 1 && N <= 6400);   // g rows staged in shared memory: N x 8 floats
  DDIMB_PROF(stream, "linear_small_dgrad", N, K, 4.0 * ((double)N * K + (double)rows * (2 * K + N)), 2.0 * rows * N * K);
  size_t smem = (size_t)N * kLsdRows * sizeof(float);
  const size_t red = (size_t)kLsdWarps * kLsdRows * 32 * sizeof(float);
  if (smem < red) smem = red;
  DDIMB_CHECK_ARG(smem <= 200 * 1024);
  DDIMB_SMEM_ATTR(linear_small_dgrad_kernel, 200 * 1024);
  linear_small_dgrad_kernel<<<dim3((K + 31) / 32, (rows + kLsdRows - 1) / kLsdRows), 32 * kLsdWarps, smem, stream>>>(g, w, z, out, rows, K, N);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ conv_in / conv_out parameter gradients
// Both boundary convolutions pair a 32-channel bf16 NHWC tensor V (a sum of two) with a 2-plane fp32 NCHW image P through a 3x3 stencil:
//   conv_in  (models/diffusion.py:189-198): dw[co][ci][ky][kx] += sum g[b,t,f,co] * x[b,ci,t+ky-1,f+kx-1], db[co] += sum g;  V = g1 + g2, P = x
//   conv_out (models/diffusion.py:199-208,284): dw[co][ci][ky][kx] += sum dout[b,co,t,f] * hs[b,t+ky-1,f+kx-1,ci], db[co] += sum dout;
//            V = h + skip, P = dout, i.e. hs[q] meets dout[q - (ky-1, kx-1)]
// so one kernel accumulates  acc[c][pl][dy][dx] = sum_q V[q][c] * P[pl][q + (dy, dx)]  and the two ops differ only in where an accumulator
// lands (conv_out mirrors the tap).  HBM-bound on reading V once (P is 1/16 of it).  A CTA takes 8 rows of one sample per unit (warp = row):
// the 10 P rows it meets sit in shared memory with zeroed halo columns; a lane owns 4 channels (one 8-byte load per position and tensor)
// and a quarter of the row, and keeps its 4 x 18 sums in registers across all units of a grid-stride loop; one reduction and one set of
// atomics per CTA at the end.
constexpr int kBwRows = 8, kBwThreads = 256;
__device__ __forceinline__ int bw_sidx(int c) { return c + ((c >> 6) << 3); }   // row segments of 64 land on different banks

template <bool CONV_OUT>
__global__ void __launch_bounds__(kBwThreads, 2) boundary_wgrad_kernel(const __nv_bfloat16* __restrict__ v1, const __nv_bfloat16* __restrict__ v2,
                                                                       const float* __restrict__ img, float* __restrict__ dw, float* __restrict__ db,
                                                                       int B, int T, int F, int units_t, int num_units, int pitch) {
  extern __shared__ float s_p[];                           // [10 rows][2 planes][pitch]; reused as [8 warps][76][8] at the end
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cg = lane & 7, seg = lane >> 3;
  const int L = (F + 3) >> 2;
  const int f_lo = seg * L, f_hi = min(F, f_lo + L);
  float acc[4][18];
#pragma unroll
  for (int c = 0; c < 4; ++c)
#pragma unroll
    for (int i = 0; i < 18; ++i) acc[c][i] = 0.f;
  float vsum[4] = {0.f, 0.f, 0.f, 0.f};
  float psum[2] = {0.f, 0.f};

  for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
    const int b = unit / units_t;
    const int t0 = (unit - b * units_t) * kBwRows;
    __syncthreads();                                       // the previous unit's readers are done with s_p
    for (int r = warp; r < kBwRows + 2; r += kBwThreads / 32) {
      const int t = t0 - 1 + r;
      const bool in_t = t >= 0 && t < T;
#pragma unroll
      for (int pl = 0; pl < 2; ++pl) {
        const float* src = img + ((int64_t)(b * 2 + pl) * T + t) * F;
        float* dst = s_p + (r * 2 + pl) * pitch;
        float own = 0.f;
        for (int c = lane; c < F + 2; c += 32) {
          const int f = c - 1;
          const float val = (in_t && f >= 0 && f < F) ? __ldg(src + f) : 0.f;
          dst[bw_sidx(c)] = val;
          own += val;
        }
        if (CONV_OUT && r >= 1 && r <= kBwRows) psum[pl] += own;
      }
    }
    __syncthreads();
    const int t = t0 + warp;
    if (t < T) {
      const int64_t rowoff = (((int64_t)b * T + t) * F) * 32 + cg * 4;
      const float* p0 = s_p + (warp * 2) * pitch;          // rows t-1, t, t+1 = warp, warp+1, warp+2
      for (int f0 = f_lo; f0 < f_hi; f0 += 4) {
        uint2 q1[4], q2[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int f = min(f0 + u, f_hi - 1);
          q1[u] = *reinterpret_cast<const uint2*>(v1 + rowoff + (int64_t)f * 32);
          if (v2) q2[u] = *reinterpret_cast<const uint2*>(v2 + rowoff + (int64_t)f * 32);
        }
        // stencil columns f0-1 .. f0+4 of the three rows and two planes
        float win[2][3][6];
#pragma unroll
        for (int pl = 0; pl < 2; ++pl)
#pragma unroll
          for (int dy = 0; dy < 3; ++dy)
#pragma unroll
            for (int j = 0; j < 6; ++j) {
              const int c = min(f0 + j, F + 1);            // halo-shifted column of f0 - 1 + j
              win[pl][dy][j] = p0[(dy * 2 + pl) * pitch + bw_sidx(c)];
            }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (f0 + u >= f_hi) break;
          float2 a = unpack_bf16x2(q1[u].x), c2 = unpack_bf16x2(q1[u].y);
          float v[4] = {a.x, a.y, c2.x, c2.y};
          if (v2) {
            a = unpack_bf16x2(q2[u].x); c2 = unpack_bf16x2(q2[u].y);
            v[0] += a.x; v[1] += a.y; v[2] += c2.x; v[3] += c2.y;
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (!CONV_OUT) vsum[c] += v[c];
#pragma unroll
            for (int pl = 0; pl < 2; ++pl)
#pragma unroll
              for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) acc[c][(pl * 3 + dy) * 3 + dx] = fmaf(v[c], win[pl][dy][u + dx], acc[c][(pl * 3 + dy) * 3 + dx]);
          }
        }
      }
    }
  }

  // reduce the four row quarters (lanes 8 apart), then the eight warps, then one atomic per accumulator
#pragma unroll
  for (int c = 0; c < 4; ++c) {
#pragma unroll
    for (int i = 0; i < 18; ++i) {
      float s = acc[c][i];
      s += __shfl_xor_sync(0xffffffffu, s, 8);
      s += __shfl_xor_sync(0xffffffffu, s, 16);
      acc[c][i] = s;
    }
    vsum[c] += __shfl_xor_sync(0xffffffffu, vsum[c], 8);
    vsum[c] += __shfl_xor_sync(0xffffffffu, vsum[c], 16);
  }
#pragma unroll
  for (int pl = 0; pl < 2; ++pl)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) psum[pl] += __shfl_xor_sync(0xffffffffu, psum[pl], o);
  __syncthreads();
  float* red = s_p;                                        // [8 warps][32 channels][19] + [8][2]
  if (seg == 0) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
#pragma unroll
      for (int i = 0; i < 18; ++i) red[(warp * 32 + cg * 4 + c) * 19 + i] = acc[c][i];
      red[(warp * 32 + cg * 4 + c) * 19 + 18] = vsum[c];
    }
  }
  if (lane == 0) { red[8 * 32 * 19 + warp * 2] = psum[0]; red[8 * 32 * 19 + warp * 2 + 1] = psum[1]; }
  __syncthreads();
  for (int o = threadIdx.x; o < 32 * 19; o += blockDim.x) {
    const int ch = o / 19, i = o - ch * 19;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[(w * 32 + ch) * 19 + i];
    if (i == 18) {
      if (!CONV_OUT) atomicAdd(db + ch, s);
      continue;
    }
    const int pl = i / 9, dy = (i - pl * 9) / 3, dx = i - pl * 9 - dy * 3;
    if (CONV_OUT) atomicAdd(dw + ((int64_t)pl * 32 + ch) * 9 + (2 - dy) * 3 + (2 - dx), s);   // hs[q] meets dout[q + d]  <=>  tap k = 1 - d
    else atomicAdd(dw + ((int64_t)ch * 2 + pl) * 9 + dy * 3 + dx, s);                       // g[q] meets x[q + d]      <=>  tap k = 1 + d
  }
  if (CONV_OUT && threadIdx.x < 2) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[8 * 32 * 19 + w * 2 + threadIdx.x];
    atomicAdd(db + threadIdx.x, s);
  }
}

template <bool CONV_OUT>
static int boundary_wgrad_launch(const void* v1, const void* v2, const float* img, float* dw, float* db, int B, int T, int F, cudaStream_t stream) {
  const int c_max = F + 1;
  const int pitch = ((c_max + ((c_max >> 6) << 3) + 1) + 7) & ~7;
  size_t smem = (size_t)(kBwRows + 2) * 2 * pitch * sizeof(float);
  const size_t red = (size_t)(8 * 32 * 19 + 16) * sizeof(float);
  if (smem < red) smem = red;
  DDIMB_SMEM_ATTR((boundary_wgrad_kernel<CONV_OUT>), smem);
  const int units_t = (T + kBwRows - 1) / kBwRows;
  const int64_t units = (int64_t)B * units_t;
  int64_t grid = (int64_t)ddimb_num_sms() * 2;
  if (grid > units) grid = units;
  boundary_wgrad_kernel<CONV_OUT><<<(unsigned)grid, kBwThreads, smem, stream>>>(reinterpret_cast<const __nv_bfloat16*>(v1), reinterpret_cast<const __nv_bfloat16*>(v2),
                                                                               img, dw, db, B, T, F, units_t, (int)units, pitch);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

extern "C" int ddimb_conv_in_wgrad(const float* x, const void* g1, const void* g2, float* dw, float* db, int B, int Cin, int T, int F,
                                   int Cout, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && g1 && dw && db && B >= 1 && T >= 1 && F >= 1 && Cin == 2 && Cout == 32);
  DDIMB_CHECK_ARG((int64_t)B * ((T + kBwRows - 1) / kBwRows) < (1ll << 31));
  DDIMB_PROF(stream, "conv_in_wgrad", Cin, Cout, (double)B * T * F * (4.0 * Cin + 2.0 * Cout * (g2 ? 2 : 1)), 18.0 * Cin * Cout * B * T * F);
  return boundary_wgrad_launch<false>(g1, g2, x, dw, db, B, T, F, stream);
}

extern "C" int ddimb_conv_out_wgrad(const void* h, const void* skip, const float* dout, float* dw, float* db, int B, int T, int F, int Cin,
                                    int Cout, cudaStream_t stream) {
  DDIMB_CHECK_ARG(h && dout && dw && db && B >= 1 && T >= 1 && F >= 1 && Cin == 32 && Cout == 2);
  DDIMB_CHECK_ARG((int64_t)B * ((T + kBwRows - 1) / kBwRows) < (1ll << 31));
  DDIMB_PROF(stream, "conv_out_wgrad", Cin, Cout, (double)B * T * F * (2.0 * Cin * (skip ? 2 : 1) + 4.0 * Cout), 18.0 * Cin * Cout * B * T * F);
  return boundary_wgrad_launch<true>(h, skip, dout, dw, db, B, T, F, stream);
}
