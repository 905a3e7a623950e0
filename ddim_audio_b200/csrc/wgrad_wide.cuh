// Weight gradient of the 3x3 convolutions of the widest level (C = 32, W a multiple of 128): autograd of reference
// models/diffusion.py:28-40,  dw[co][ci][ky][kx] += sum_{b,t,w} dy[b,t,w,co] * x[b,t+ky-1,w+kx-1,ci].
//
// The contraction runs over 10^8 positions against a 9 x 32 x 32 result: the op is HBM-bound on reading x and dy ONCE (arithmetic
// intensity 144 flop/B).  The generic split-K kernel (gemm_tn.cuh) re-reads both tensors for each of the 9 taps (r02: 0.095 of
// HBM, 17 % of the train step).  Here a CTA walks a 128-column strip down consecutive rows like the forward kernel
// (conv_wide.cuh): every x row strip (144 positions: halo + padding to a multiple of 16, TMA zero fill outside the image) and
// every dy row strip (128 positions between zeroed guard rows) is read once into a ring slot, and ALL NINE accumulators
// [ci][co] stay in tensor memory for the CTA's whole lifetime; they are added to dw once at the end.
//   * both operands are MN-major (channels contiguous per position, exactly as TMA delivers NHWC rows); K = positions.
//   * the three kx taps of one ky are ONE tcgen05.mma of N = 96: the B descriptor's leading-dimension offset is one position row
//     (64 B), so its three 32-channel atoms are the dy tile shifted by 0, 1, 2 positions (atom j <-> kx = 2 - j).
//   * ky = which x row meets the dy row: x row r feeds (dy row r - 2, ky = 2), (r - 1, ky = 1), (r, ky = 0).
// Warp roles (192 threads): 0 = TMA producer, 1 = MMA issuer, 4 = final reduction (TMEM lane quarter 0: lanes 0..31 = ci), others idle.
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kWgThreads = 192;
constexpr int kWgC = 32;
constexpr int kWgXRows = 144;                          // x positions per slot: w0 - 1 .. w0 + 142
constexpr int kWgXSlot = kWgXRows * 64;                // 9216
constexpr int kWgYLead = 8;                            // zeroed guard rows ahead of the 128 dy positions (keeps the TMA destination 512-B aligned)
constexpr int kWgYRows = 152;                          // 8 guard + 128 data + 16 guard
constexpr int kWgYSlot = ((kWgYRows * 64 + 1023) / 1024) * 1024;   // 10240
constexpr int kWgXSlots = 8, kWgYSlots = 6;
constexpr int kWgSmem = 1024 + kWgXSlots * kWgXSlot + kWgYSlots * kWgYSlot + 1024;

struct WgParams {
  int32_t B, T, W;
  int32_t RT, units_t, num_units;
  float* dw;              // fp32 [32][32][3][3], accumulated
};

struct WgUnit {
  int b, w0, t0, rows;
  __device__ __forceinline__ void set(const WgParams& p, int u) {
    const int per_b = (p.W >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    w0 = wb << 7;
    rows = min(p.RT, p.T - t0);
  }
};

__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_wide_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmY, const __grid_constant__ WgParams p) {
  constexpr int NX = kWgXSlots, NY = kWgYSlots, C = kWgC;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t x_base = smem_base, y_base = smem_base + NX * kWgXSlot;
  uint8_t* y_gen = smem_gen + NX * kWgXSlot;
  const uint32_t aux = y_base + NY * kWgYSlot;
  uint8_t* aux_gen = y_gen + NY * kWgYSlot;
  // aux: x_full[8] @0, x_empty[8] @64, y_full[8] @128, y_empty[8] @192, done @256, tmem ptr @264
  const uint32_t bar_xf = aux, bar_xe = aux + 64, bar_yf = aux + 128, bar_ye = aux + 192, bar_done = aux + 256;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 264);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // guard rows of every dy slot are zero for the kernel's lifetime (TMA only ever writes the 128 data rows)
  for (int i = threadIdx.x; i < NY * kWgYSlot / 16; i += blockDim.x) reinterpret_cast<uint4*>(y_gen)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmY);
    for (int s = 0; s < NX; ++s) { mbar_init(bar_xf + 8 * s, 1); mbar_init(bar_xe + 8 * s, 1); }
    for (int s = 0; s < NY; ++s) { mbar_init(bar_yf + 8 * s, 1); mbar_init(bar_ye + 8 * s, 1); }
    mbar_init(bar_done, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 264, 512u);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ TMA producer: per unit x rows t0-1 .. t0+rows, dy rows t0 .. t0+rows-1
    if (elect_one()) {
      uint32_t xc = 0, yc = 0;
      WgUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr) {
          {
            const uint32_t slot = xc % NX, use = xc / NX;
            mbar_wait(bar_xe + 8 * slot, (use & 1) ^ 1, 81);
            mbar_arrive_expect_tx(bar_xf + 8 * slot, (uint32_t)kWgXSlot);
            tma_load_4d(x_base + slot * kWgXSlot, &tmX, bar_xf + 8 * slot, 0, u.w0 - 1, u.t0 - 1 + rr, u.b);
            ++xc;
          }
          if (rr < u.rows) {
            const uint32_t slot = yc % NY, use = yc / NY;
            mbar_wait(bar_ye + 8 * slot, (use & 1) ^ 1, 82);
            mbar_arrive_expect_tx(bar_yf + 8 * slot, (uint32_t)(128 * 64));
            tma_load_4d(y_base + slot * kWgYSlot + kWgYLead * 64, &tmY, bar_yf + 8 * slot, 0, u.w0, u.t0 + rr, u.b);
            ++yc;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread)
    if (elect_one()) {
      // MN-major SWIZZLE_64B operands: atom = [8 positions][32 channels]; SBO = 8 position rows.  A: M = 128 rows of D, of which only
      // the first 32-channel atom exists -- its three phantom atoms (LBO = one row) produce D rows 32..127, which are never read.
      // B: N = 96 = three atoms one position row apart.
      const uint64_t hi = ((uint64_t)(64 >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)4 << 61);
      const uint32_t idesc = make_idesc_bf16(128, 96) | (1u << 15) | (1u << 16);
      uint32_t xc = 0, yc = 0;
      uint32_t first = 7u;                                   // accumulators ky = 0, 1, 2 not written yet
      WgUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr) {
          const uint32_t xs = (xc + rr) % NX;
          mbar_wait(bar_xf + 8 * xs, ((xc + rr) / NX) & 1, 83);
          if (rr < u.rows) mbar_wait(bar_yf + 8 * ((yc + rr) % NY), ((yc + rr) / NY) & 1, 84);   // dy row rr is first needed now (ky = 0)
          tc_fence_after();
          const int t = u.t0 - 1 + rr;
          if (t >= 0 && t < p.T) {                           // an x row outside the image is all zero padding
            const uint32_t a_lo = ((x_base + xs * kWgXSlot) & 0x3FFFFu) >> 4;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
              const int r = rr - ky;                         // dy row met through kernel row ky
              if (r < 0 || r >= u.rows) continue;
              const uint32_t ys = (yc + (uint32_t)r) % NY;
              const uint32_t b_lo = ((y_base + ys * kWgYSlot + (kWgYLead - 2) * 64) & 0x3FFFFu) >> 4;   // row 6: position w0 - 2
              const uint32_t acc = tmem_base + (uint32_t)(ky * 96);
              const uint32_t f = (first >> ky) & 1u;
#pragma unroll
              for (int s = 0; s < kWgXRows / 16; ++s)
                umma_bf16(acc, hi | (uint64_t)(a_lo + (uint32_t)(s * 64)), hi | (uint64_t)(b_lo + (uint32_t)(s * 64)), idesc, (s == 0 && f) ? 0u : 1u);
              first &= ~(1u << ky);
            }
          }
          umma_commit(bar_xe + 8 * xs);                      // x row consumed
          if (rr >= 2) umma_commit(bar_ye + 8 * ((yc + (uint32_t)(rr - 2)) % NY));   // dy row rr - 2 has met its three x rows
        }
        xc += (uint32_t)(u.rows + 2);
        yc += (uint32_t)u.rows;
      }
      *reinterpret_cast<volatile uint32_t*>(aux_gen + 272) = first;   // accumulators that were never written (T = 1: no ky = 0 / 2 pair exists)
      __threadfence_block();
      umma_commit(bar_done);
    }
  } else if (warp == 4) {
    // ================================================================ final reduction: lanes 0..31 of TMEM = ci; column ky * 96 + j * 32 + co, kx = 2 - j
    mbar_wait(bar_done, 0, 85);
    tc_fence_after();
    const uint32_t unwritten = *reinterpret_cast<volatile uint32_t*>(aux_gen + 272);
#pragma unroll 1
    for (int c0 = 0; c0 < 288; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + (uint32_t)c0, v);
      tmem_ld_wait();
      const int ky = c0 / 96;
      if (!((unwritten >> ky) & 1u)) {
        const int rem = c0 - ky * 96, kx = 2 - rem / 32, co0 = rem & 31;
#pragma unroll
        for (int i = 0; i < 16; ++i) atomicAdd(p.dw + ((int64_t)(co0 + i) * C + lane) * 9 + ky * 3 + kx, __uint_as_float(v[i]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace ddimb
