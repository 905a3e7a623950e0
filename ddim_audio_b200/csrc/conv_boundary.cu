// The two convolutions at the network boundary (reference models/diffusion.py:189-208): conv_in 2 -> 32 on the fp32 NCHW input
// and conv_out 32 -> 2 of (x + skip) back to fp32 NCHW.  K = 18 / N = 2 are not tensor-core shapes; both are HBM-bound
// (66-132 B per position) with 576 FMAs per position, so the kernels are built around the FMA pipe:
//   * the 576-float weight sits in a __constant__ bank (staged device-to-device on the caller's stream), so every FMA takes
//     its weight as a constant operand -- no shared-memory or register traffic for weights;
//   * a CTA walks a 128-column strip down RT rows: conv_in keeps a 3x3x2 input window in registers, conv_out keeps a
//     3-row fp32 ring of (x + skip) in shared memory (each input row read from HBM once, halo columns included);
//   * conv_in's GroupNorm partial sums stay in registers for the whole strip (one set of fp64 atomics per CTA).
// The constant bank is process-global: calls on DIFFERENT streams with different weights must not overlap.
#include <stdlib.h>
#include "../../include/ddim_b200.h"
#include "common.cuh"

using namespace ddimb;

namespace {

constexpr int kStripRows = 32;

__constant__ float c_w_in[32 * 2 * 9 + 32];    // [co][ci][ky][kx] | bias[co]
__constant__ float c_w_out[2 * 32 * 9 + 2];    // [co][ci][ky][kx] | bias[co]

// ------------------------------------------------------------------------------------------------ conv_in
template <int CIN, int CO>
__global__ void __launch_bounds__(128) conv_in_strip_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                           double* __restrict__ stats, int T, int F, int strips) {
  __shared__ float s_red[4][16];
  const int unit = blockIdx.x;
  const int strip = unit % strips;
  const int t0 = (unit / strips) * kStripRows;
  const int b = blockIdx.y;
  const int f = strip * 128 + threadIdx.x;
  const bool active = f < F;
  const int rows = min(kStripRows, T - t0);
  float win[CIN][3][3];   // [ci][row t-1, t, t+1][col f-1, f, f+1]
  auto load_row = [&](int tt, float (&dst)[CIN][3][3], int slot) {
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci) {
      const float* xp = x + ((int64_t)(b * CIN + ci) * T + tt) * F;
      const bool ok = active && tt >= 0 && tt < T;
      dst[ci][slot][0] = (ok && f >= 1) ? __ldg(xp + f - 1) : 0.f;
      dst[ci][slot][1] = ok ? __ldg(xp + f) : 0.f;
      dst[ci][slot][2] = (ok && f + 1 < F) ? __ldg(xp + f + 1) : 0.f;
    }
  };
  load_row(t0 - 1, win, 0);
  load_row(t0, win, 1);
  float sacc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) sacc[k] = 0.f;
  constexpr int CG = CO / 8;
  for (int r = 0; r < rows; ++r) {
    const int t = t0 + r;
    load_row(t + 1, win, 2);
    float acc[CO];
#pragma unroll
    for (int c = 0; c < CO; ++c) acc[c] = c_w_in[CO * CIN * 9 + c];
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float v = win[ci][ky][kx];
#pragma unroll
          for (int c = 0; c < CO; ++c) acc[c] = fmaf(v, c_w_in[(c * CIN + ci) * 9 + ky * 3 + kx], acc[c]);
        }
    if (active) {
      uint4* op = reinterpret_cast<uint4*>(out + (((int64_t)b * T + t) * F + f) * CO);
#pragma unroll
      for (int c = 0; c < CO; c += 8)
        stg_v4(op + (c >> 3), make_uint4(pack_bf16x2(acc[c], acc[c + 1]), pack_bf16x2(acc[c + 2], acc[c + 3]),
                                         pack_bf16x2(acc[c + 4], acc[c + 5]), pack_bf16x2(acc[c + 6], acc[c + 7])));
      if (stats) {
#pragma unroll
        for (int g = 0; g < 8; ++g)
#pragma unroll
          for (int j = 0; j < CG; ++j) {
            const float v = acc[g * CG + j];
            sacc[2 * g] += v;
            sacc[2 * g + 1] = fmaf(v, v, sacc[2 * g + 1]);
          }
      }
    }
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) { win[ci][0][kx] = win[ci][1][kx]; win[ci][1][kx] = win[ci][2][kx]; }
  }
  if (stats) {
    // fixed-order reduction: lanes by shuffle tree, the four warps in warp order, then one fp64 atomic per value per CTA
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float v = warp_sum(sacc[k]);
      if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 16)
      atomicAdd(stats + b * 16 + threadIdx.x,
                (double)((s_red[0][threadIdx.x] + s_red[1][threadIdx.x]) + (s_red[2][threadIdx.x] + s_red[3][threadIdx.x])));
  }
}

// ------------------------------------------------------------------------------------------------ conv_out
// ring: 3 rows x 130 positions x 32 fp32 channels; float4 j of position p lives at chunk j ^ (p & 7) (bank-conflict-free for a
// warp reading the same j at consecutive p)
constexpr int kRingPos = 130;
constexpr int kRingRowFloats = kRingPos * 32;

template <int CI, int CO>
__global__ void __launch_bounds__(128, 3) conv_out_strip_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ x2,
                                                            float* __restrict__ out, int T, int F, int strips) {
  extern __shared__ float4 ring4[];   // [3][130][8] float4
  const int unit = blockIdx.x;
  const int strip = unit % strips;
  const int t0 = (unit / strips) * kStripRows;
  const int b = blockIdx.y;
  const int w0 = strip * 128;
  const int rows = min(kStripRows, T - t0);
  constexpr int ITEMS = kRingPos * (CI / 8);                 // 16-byte bf16 chunks per row strip
  constexpr int NIT = (ITEMS + 127) / 128;
  uint4 pre[NIT], pre2[NIT];
  auto fetch = [&](int tt) {
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
      const int item = threadIdx.x + i * 128;
      const int p = item / (CI / 8), ch = item % (CI / 8);
      const int col = w0 - 1 + p;
      pre[i] = make_uint4(0u, 0u, 0u, 0u);
      pre2[i] = make_uint4(0u, 0u, 0u, 0u);
      if (item < ITEMS && tt >= 0 && tt < T && col >= 0 && col < F) {
        const int64_t off = (((int64_t)b * T + tt) * F + col) * CI + ch * 8;
        pre[i] = ldg_nc_v4(x + off);
        if (x2) pre2[i] = ldg_nc_v4(x2 + off);
      }
    }
  };
  auto stash = [&](int slot) {
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
      const int item = threadIdx.x + i * 128;
      if (item < ITEMS) {
        const int p = item / (CI / 8), ch = item % (CI / 8);
        const uint32_t a[4] = {pre[i].x, pre[i].y, pre[i].z, pre[i].w}, c[4] = {pre2[i].x, pre2[i].y, pre2[i].z, pre2[i].w};
        float v[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 u = unpack_bf16x2(a[j]), w = unpack_bf16x2(c[j]);   // the additive skip of models/diffusion.py:284
          v[2 * j] = u.x + w.x; v[2 * j + 1] = u.y + w.y;
        }
        float4* dst = ring4 + (slot * kRingPos + p) * 8;
        dst[(2 * ch) ^ (p & 7)] = make_float4(v[0], v[1], v[2], v[3]);
        dst[(2 * ch + 1) ^ (p & 7)] = make_float4(v[4], v[5], v[6], v[7]);
      }
    }
  };
  // rows t0-1, t0, t0+1 into slots 0, 1, 2; inside the loop row t+2 is fetched into registers while row t is computed
  fetch(t0 - 1); stash(0);
  fetch(t0); stash(1);
  fetch(t0 + 1); stash(2);
  __syncthreads();
  const int f = w0 + threadIdx.x;
  for (int r = 0; r < rows; ++r) {
    const int t = t0 + r;
    if (r + 1 < rows) fetch(t + 2);   // global loads in flight during the FMA block below
    float acc[CO][4];
#pragma unroll
    for (int co = 0; co < CO; ++co) { acc[co][0] = c_w_out[CO * CI * 9 + co]; acc[co][1] = acc[co][2] = acc[co][3] = 0.f; }
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int slot = (r + ky) % 3;      // row t-1+ky
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int p = threadIdx.x + kx;   // ring position of column f-1+kx
        const float4* src = ring4 + (slot * kRingPos + p) * 8;
#pragma unroll
        for (int j = 0; j < CI / 4; ++j) {
          const float4 v = src[j ^ (p & 7)];
#pragma unroll
          for (int co = 0; co < CO; ++co) {
            const int wb = (co * CI + 4 * j) * 9 + ky * 3 + kx;
            acc[co][0] = fmaf(v.x, c_w_out[wb], acc[co][0]);
            acc[co][1] = fmaf(v.y, c_w_out[wb + 9], acc[co][1]);
            acc[co][2] = fmaf(v.z, c_w_out[wb + 18], acc[co][2]);
            acc[co][3] = fmaf(v.w, c_w_out[wb + 27], acc[co][3]);
          }
        }
      }
    }
    if (f < F) {
#pragma unroll
      for (int co = 0; co < CO; ++co)
        out[(((int64_t)b * CO + co) * T + t) * F + f] = (acc[co][0] + acc[co][1]) + (acc[co][2] + acc[co][3]);
    }
    if (r + 1 < rows) {
      __syncthreads();              // every thread is done with the slot of row t-1 ...
      stash(r % 3);                 // ... which now receives row t+2
      __syncthreads();
    }
  }
}

}  // namespace

int ddimb_conv_in_tc(const float* x, const uint8_t* wimg, const float* bias, void* out, double* stats, int B, int T, int F, cudaStream_t stream);

// models/diffusion.py:189-198 -- Conv2d(Cin=2 -> 32, k3 p1): x fp32 NCHW, w fp32 [32][Cin][3][3], out bf16 NHWC, GroupNorm partial sums.
extern "C" int ddimb_conv_in(const float* x, const float* w, const void* wimg, const float* bias, void* out, double* stats, int B, int Cin, int T,
                             int F, int Cout, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && w && bias && out && B >= 1 && Cin == 2 && Cout == 32 && B <= 65535);
  DDIMB_PROF(stream, "conv_in", Cin, Cout, (double)B * T * F * (4.0 * Cin + 2.0 * Cout), 18.0 * Cin * Cout * B * T * F);
  if (F % 128 == 0)   // tensor-core path (prologue 3 of csrc/conv_wide.cuh); the strip kernel below covers ragged F
    return ddimb_conv_in_tc(x, reinterpret_cast<const uint8_t*>(wimg), bias, out, stats, B, T, F, stream);
  DDIMB_CUDA(cudaMemcpyToSymbolAsync(c_w_in, w, sizeof(float) * 32 * 2 * 9, 0, cudaMemcpyDeviceToDevice, stream));
  DDIMB_CUDA(cudaMemcpyToSymbolAsync(c_w_in, bias, sizeof(float) * 32, sizeof(float) * 32 * 2 * 9, cudaMemcpyDeviceToDevice, stream));
  const int strips = (F + 127) / 128;
  const int units = strips * ((T + kStripRows - 1) / kStripRows);
  conv_in_strip_kernel<2, 32><<<dim3(units, B), 128, 0, stream>>>(x, reinterpret_cast<__nv_bfloat16*>(out), stats, T, F, strips);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

int ddimb_conv_out_tc(const void* x, const void* skip, const uint8_t* wimg, const float* bias, float* out, int B, int T, int F, cudaStream_t stream);

// models/diffusion.py:284,199-208 -- Conv2d(32 -> 2, k3 p1) of (x + skip): x, skip (nullable) bf16 NHWC, w fp32 [2][32][3][3], out fp32 NCHW.
extern "C" int ddimb_conv_out(const void* x, const void* skip, const float* w, const void* wimg, const float* bias, float* out, int B, int T, int F,
                              int Cin, int Cout, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && w && bias && out && B >= 1 && Cin == 32 && Cout == 2 && B <= 65535);
  DDIMB_PROF(stream, "conv_out", Cin, Cout, (double)B * T * F * (2.0 * Cin * (skip ? 2 : 1) + 4.0 * Cout), 18.0 * Cin * Cout * B * T * F);
  if (F % 128 == 0)   // tensor-core path (csrc/conv_out_tc.cuh); the strip kernel below covers ragged F
    return ddimb_conv_out_tc(x, skip, reinterpret_cast<const uint8_t*>(wimg), bias, out, B, T, F, stream);
  DDIMB_CUDA(cudaMemcpyToSymbolAsync(c_w_out, w, sizeof(float) * 2 * 32 * 9, 0, cudaMemcpyDeviceToDevice, stream));
  DDIMB_CUDA(cudaMemcpyToSymbolAsync(c_w_out, bias, sizeof(float) * 2, sizeof(float) * 2 * 32 * 9, cudaMemcpyDeviceToDevice, stream));
  const int strips = (F + 127) / 128;
  const int units = strips * ((T + kStripRows - 1) / kStripRows);
  const int smem = 3 * kRingRowFloats * (int)sizeof(float);
  auto strip_kernel = conv_out_strip_kernel<32, 2>;
  DDIMB_SMEM_ATTR(strip_kernel, smem);
  conv_out_strip_kernel<32, 2><<<dim3(units, B), 128, smem, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                                      reinterpret_cast<const __nv_bfloat16*>(skip), out, T, F, strips);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}
