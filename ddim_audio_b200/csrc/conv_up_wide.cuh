// Upsample ConvTranspose2d(64 -> 32, k4 s2 p1) + additive skip into the widest level (reference models/diffusion.py:59-67,284),
// where the generic kernel is bound by TMA box rows (6 activation + 8 weight boxes per 128 inputs and output-row parity).
// Same walk as csrc/conv_wide.cuh: a CTA takes a 128-column strip of INPUT rows top to bottom, every input row is read from
// HBM once into a ring slot (130 positions x 128 B, halo zero-filled by TMA), the weights (96 KB) stay resident.
//   out[2t+r, 2f+s, co] = sum_{dy,dx} in[t+dy, f+dx, :] . Wt[:, co, ky = r+1-2dy, kx = s+1-2dx]
//   * input row ti feeds the FOUR consecutive output rows o = 2 ti - 1 + ky (ky = 0..3); their accumulator blocks (64 columns
//     = [s][co], i.e. the 128 contiguous bytes of output pixel pair f) are adjacent in TMEM, so one MMA of N = 256 serves all.
//   * dx = -1, 0, +1 are A descriptors starting 0, 1, 2 rows into the slot; each has its own weight tile [ky][s][co] x 64 ci
//     whose rows are zero where (s, dx) has no kernel column (dx = -1: s = 1; dx = +1: s = 0).
//   * blocks ky = 0,1 complete with this input row, blocks ky = 2,3 receive their first contribution from it.
// Warp roles (320 threads): 0 = TMA producer, 1 = MMA issuer, 2-9 = epilogue (two warps per TMEM lane quarter, alternating
// output rows): bias, + skip, GroupNorm partial sums, bf16 NHWC store.
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kUwThreads = 320;
constexpr int kUwCin = 64, kUwCout = 32;
constexpr int kUwSlotBytes = ((kWidePos * 128 + 1023) / 1024) * 1024;   // 17408
constexpr int kUwSlots = 6;
constexpr int kUwNB = 8;                               // output rows in flight in TMEM (64 columns each)
constexpr int kUwWTile = 256 * 128;                    // [4 ky][2 s][32 co] rows x 64 ci
constexpr int kUwWBytes = 3 * kUwWTile;
constexpr int kUwAux = 1024 + 4096 + 8 * 2048;         // barriers | 8 bias copies | 8 x store staging
constexpr int kUwSmem = 1024 + kUwWBytes + kUwSlots * kUwSlotBytes + kUwAux;

struct UwParams {
  int32_t B, T, W;        // INPUT extents (W a multiple of 128)
  int32_t RT, units_t, num_units;
  const uint8_t* wimg;
  const float* bias;
  const __nv_bfloat16* skip;   // [B][2T][2W][32] or null
  __nv_bfloat16* out;          // [B][2T][2W][32]
  double* stats;
};

// ABI weight layout w[(ky * 4 + kx) * 32 + co][ci] (bf16, 64 ci) -> three SWIZZLE_128B tiles (dx = -1, 0, +1) of 256 rows
// n = ky * 64 + s * 32 + co; kx = s + 1 - 2 dx, rows with kx outside 0..3 are zero.
__global__ void __launch_bounds__(256) uw_weight_image_kernel(const __nv_bfloat16* __restrict__ w, uint8_t* __restrict__ dst) {
  const int tile = blockIdx.x, dx = tile - 1;
  for (int i = threadIdx.x; i < 256 * 8; i += blockDim.x) {
    const int row = i >> 3, j = i & 7;
    const int ky = row >> 6, s = (row >> 5) & 1, co = row & 31;
    const int kx = s + 1 - 2 * dx;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (kx >= 0 && kx <= 3) v = *reinterpret_cast<const uint4*>(w + ((int64_t)(ky * 4 + kx) * kUwCout + co) * kUwCin + j * 8);
    *reinterpret_cast<uint4*>(dst + (int64_t)tile * kUwWTile + row * 128 + ((j ^ (row & 7)) << 4)) = v;
  }
}

struct UwUnit {
  int b, w0, t0, rows;   // input coordinates
  __device__ __forceinline__ void set(const UwParams& p, int u) {
    const int per_b = (p.W >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    w0 = wb << 7;
    rows = min(p.RT, p.T - t0);
  }
};

__global__ void __launch_bounds__(kUwThreads, 1)
conv_up_wide_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ UwParams p) {
  constexpr int NS = kUwSlots, NB = kUwNB, NA = 2 * kUwCout;   // accumulator block width
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t w_base = smem_base;
  const uint32_t a_base = smem_base + kUwWBytes;
  const uint32_t aux = a_base + NS * kUwSlotBytes;
  uint8_t* aux_gen = smem_gen + kUwWBytes + NS * kUwSlotBytes;
  // aux: full[8] @0, empty[8] @64, acc_full[8] @128, acc_empty[8] @192, wfull @256, tmem ptr @264
  const uint32_t bar_full = aux, bar_empty = aux + 64, bar_afull = aux + 128, bar_aempty = aux + 192, bar_w = aux + 256;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 264);
  float* bias_smem = reinterpret_cast<float*>(aux_gen + 1024);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int b = 0; b < NB; ++b) {
      mbar_init(bar_afull + 8 * b, 1);
      mbar_init(bar_aempty + 8 * b, 4);
    }
    mbar_init(bar_w, 1);
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

  // Output row with unit-local index lo (0 .. 2 rows - 1) of the unit whose first output row has running index g0 lives in
  // ring position (g0 + lo + 1) % NB: g0 is even, so the pairs (ky 0,1) and (ky 2,3) never straddle the ring end.
  if (warp == 0) {
    // ================================================================ TMA producer: input rows t0 - 1 .. t0 + rows
    if (elect_one()) {
      mbar_arrive_expect_tx(bar_w, (uint32_t)kUwWBytes);
      bulk_load_g2s(w_base, p.wimg, (uint32_t)kUwWBytes, bar_w);
      uint32_t cnt = 0;
      UwUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr, ++cnt) {
          const uint32_t slot = cnt % NS, use = cnt / NS;
          mbar_wait(bar_empty + 8 * slot, (use & 1) ^ 1, 61);
          mbar_arrive_expect_tx(bar_full + 8 * slot, (uint32_t)(kWidePos * 128));
          tma_load_4d(a_base + slot * kUwSlotBytes, &tmA, bar_full + 8 * slot, 0, u.w0 - 1, u.t0 - 1 + rr, u.b);
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread)
    if (elect_one()) {
      mbar_wait(bar_w, 0, 62);
      tc_fence_after();
      mbar_arrive(bar_afull + 0);   // ring position 0 is first used by the 8th output row (use count 1): retire its phase 0 up front
      const uint64_t desc_hi = make_kmajor_desc(0, 1024, 2);             // SWIZZLE_128B
      const uint32_t w_lo = (w_base & 0x3FFFFu) >> 4;
      constexpr uint32_t kBlk16 = (64 * 128) >> 4;                       // one ky block (64 rows) of a weight tile
      uint32_t cnt = 0, g0 = 0;
      UwUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        const int nout = 2 * u.rows;
        for (int rr = 0; rr < u.rows + 2; ++rr) {
          const uint32_t slot = (cnt + rr) % NS;
          const int lo0 = 2 * rr - 3;                                    // local output row of ky = 0
          // segment A = valid blocks among ky 0,1 (accumulate, complete after this row); segment B = among ky 2,3 (first contribution)
          const int fa = (lo0 >= 0 && lo0 < nout) ? 0 : 1, la = (lo0 + 1 >= 0 && lo0 + 1 < nout) ? 1 : 0;
          const int fb = (lo0 + 2 >= 0 && lo0 + 2 < nout) ? 2 : 3, lb = (lo0 + 3 >= 0 && lo0 + 3 < nout) ? 3 : 2;
          const bool has_a = fa <= la, has_b = fb <= lb;
          const uint32_t ra = (g0 + (uint32_t)(lo0 + fa + 1 + NB)) % NB, rb = (g0 + (uint32_t)(lo0 + fb + 1 + NB)) % NB;
          const bool merged = has_a && has_b && la == 1 && fb == 2 && rb == ra + (uint32_t)(2 - fa);
          mbar_wait(bar_full + 8 * slot, ((cnt + rr) / NS) & 1, 63);
          if (has_b) {
            for (int ky = fb; ky <= lb; ++ky) {
              const uint32_t g = g0 + (uint32_t)(lo0 + ky) + 1;          // ring position / use count of that output row
              mbar_wait(bar_aempty + 8 * (g % NB), ((g / NB) & 1) ^ 1, 64);
            }
          }
          tc_fence_after();
          const uint32_t a_lo = ((a_base + slot * kUwSlotBytes) & 0x3FFFFu) >> 4;
          const uint32_t ta = tmem_base + ra * NA, tb = tmem_base + rb * NA;
          const uint32_t ia = make_idesc_bf16(kTileM, (uint32_t)((la - fa + 1) * NA)), ib = make_idesc_bf16(kTileM, (uint32_t)((lb - fb + 1) * NA));
          const uint32_t iab = make_idesc_bf16(kTileM, (uint32_t)((la - fa + 1 + lb - fb + 1) * NA));
#pragma unroll
          for (int d = 0; d < 3; ++d) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t ad = desc_hi | (uint64_t)(a_lo + (uint32_t)(d * 8 + 2 * k));
              const uint32_t bl = w_lo + (uint32_t)d * (kUwWTile >> 4) + 2 * k;
              if (d == 0 && k == 0) {
                if (has_a) umma_bf16(ta, ad, desc_hi | (uint64_t)(bl + (uint32_t)fa * kBlk16), ia, 1u);
                if (has_b) umma_bf16(tb, ad, desc_hi | (uint64_t)(bl + (uint32_t)fb * kBlk16), ib, 0u);
              } else if (merged) {
                umma_bf16(ta, ad, desc_hi | (uint64_t)(bl + (uint32_t)fa * kBlk16), iab, 1u);
              } else {
                if (has_a) umma_bf16(ta, ad, desc_hi | (uint64_t)(bl + (uint32_t)fa * kBlk16), ia, 1u);
                if (has_b) umma_bf16(tb, ad, desc_hi | (uint64_t)(bl + (uint32_t)fb * kBlk16), ib, 1u);
              }
            }
          }
          umma_commit(bar_empty + 8 * slot);
          if (has_a) {
            for (int ky = fa; ky <= la; ++ky) umma_commit(bar_afull + 8 * ((g0 + (uint32_t)(lo0 + ky) + 1) % NB));
          }
        }
        cnt += u.rows + 2;
        g0 += (uint32_t)nout;
      }
    }
  } else {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4; group = output row parity)
    const int q = warp & 3;
    const uint32_t grp = (uint32_t)(warp - 2) >> 2;
    const int m = q * 32 + lane;
    float* wb = bias_smem + (warp - 2) * 128;
    uint8_t* stage = aux_gen + 1024 + 4096 + (warp - 2) * 2048;
    ConvEpiState<kUwCout> st;
    st.reset();
    st.cur_batch = -1;
    if (grp == 0 && lane == 0) mbar_arrive(bar_aempty + 0);   // same for the "accumulator free" barrier of ring position 0 (4 arrivals)
    const int To = 2 * p.T, Wo = 2 * p.W;
    uint32_t g0 = 0;
    UwUnit u;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      u.set(p, unit);
      if (u.b != st.cur_batch) {
        st.flush(p.stats, lane);
        st.cur_batch = u.b;
        conv_epilogue_bias<kUwCout>(wb, p.bias, nullptr, u.b, 0, 0, lane);
      }
      const int nout = 2 * u.rows;
      // the skip rows do not depend on the MMAs: this tile's words were requested one tile ago, the next tile's are requested
      // before waiting for the accumulator (two batches of 8 x 16 B per thread in flight)
      const int lo_first = (int)((grp - g0) & 1u);
      const int64_t base0 = ((((int64_t)u.b * To + 2 * u.t0) * Wo) + 2 * (u.w0 + m)) * kUwCout;
      const int64_t row_pitch = (int64_t)Wo * kUwCout;
      uint4 cur[NA / 8], nxt[NA / 8];
      conv_epilogue_prefetch_add<NA>(p.skip, base0 + lo_first * row_pitch, lo_first < nout, cur);
      for (int lo = lo_first; lo < nout; lo += 2) {
        const uint32_t g = g0 + (uint32_t)lo + 1;
        const uint32_t buf = g % NB, use = g / NB;
        const int64_t addr0 = base0 + lo * row_pitch;
        conv_epilogue_prefetch_add<NA>(p.skip, addr0 + 2 * row_pitch, lo + 2 < nout, nxt);
        mbar_wait(bar_afull + 8 * buf, use & 1, 65);
        tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * NA;
        conv_epilogue_tile<NA, kUwCout>(trow, wb, 0, p.skip, p.out, addr0, true, p.stats != nullptr, st, stage, p.skip != nullptr ? cur : nullptr);
#pragma unroll
        for (int k = 0; k < NA / 8; ++k) cur[k] = nxt[k];
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
      }
      g0 += (uint32_t)nout;
    }
    st.flush(p.stats, lane);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace ddimb
