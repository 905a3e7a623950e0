// Downsample Conv2d(32 -> 64, k4 s2 p1) of the widest level (reference models/diffusion.py:70-78) where the generic kernel is
// bound by TMA box rows (16 activation boxes + 16 weight boxes of 64-byte rows per 128 outputs).  Same walk as
// csrc/conv_wide.cuh: a CTA produces a 128-column strip of output rows top to bottom, every INPUT row is read from HBM once
// into a ring slot and the weights (64 KB) stay resident in shared memory.
//   * The NHWC input row is viewed space-to-depth: pixel pair f' = (2f', 2f'+1) is one 128-byte row of 64 channels, so a slot
//     is 130 pair-positions x 128 B (SWIZZLE_128B, halo pairs -1 and 128 zero-filled by TMA).
//   * kernel columns: kx = 1,2 are the two halves of pair f' (one K = 64 group, A start shifted by one row); kx = 0 is the odd
//     half of pair f'-1 (K = 32 at byte 64 of the unshifted rows); kx = 3 the even half of pair f'+1 (K = 32, shift two rows).
//   * kernel rows: input row r feeds output rows (r+1)/2 [ky = 0 | 1] and (r-1)/2 [ky = 2 | 3]: two adjacent accumulator blocks,
//     so one MMA of N = 128 serves both (weight tile rows = [ky of the earlier output row | ky of the later one]).
// Warp roles (192 threads): 0 = TMA producer, 1 = MMA issuer, 2-5 = epilogue (bias, GroupNorm partial sums, bf16 NHWC store).
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kDwThreads = 192;
constexpr int kDwCin = 32, kDwCout = 64;
constexpr int kDwSlotBytes = ((kWidePos * 128 + 1023) / 1024) * 1024;   // 17408
constexpr int kDwSlots = 8;
constexpr int kDwNB = 8;                               // output rows in flight in TMEM (64 columns each)
constexpr int kDwWTile = 128 * 128;                    // [2 blocks x 64 co][64 k] bf16
constexpr int kDwWBytes = 4 * kDwWTile;                // (input-row parity) x (centre | sides)
constexpr int kDwAux = 1024 + 4096 + 4 * 2048;         // barriers | bias copies | store staging
constexpr int kDwSmem = 1024 + kDwWBytes + kDwSlots * kDwSlotBytes + kDwAux;

struct DwParams {
  int32_t B, T, W;        // INPUT extents (T, W even; W / 2 a multiple of 128)
  int32_t RT, units_t, num_units;
  const uint8_t* wimg;    // pre-swizzled weight image (kDwWBytes)
  const float* bias;
  __nv_bfloat16* out;     // [B][T/2][W/2][64]
  double* stats;
};

// ABI weight layout w[co][(ky * 4 + kx) * 32 + ci] (bf16) -> four SWIZZLE_128B tiles [128 rows][64 k]:
//   tile = par * 2 + grp; par = input-row parity class (0: r odd -> blocks (ky 2, ky 0); 1: r even -> (ky 3, ky 1));
//   grp 0: k < 32 -> kx 1, k >= 32 -> kx 2;  grp 1: k < 32 -> kx 0... stored at k >= 32 (it multiplies the ODD half of pair f'-1),
//   and kx 3 at k < 32 (the EVEN half of pair f'+1).
__global__ void __launch_bounds__(256) dw_weight_image_kernel(const __nv_bfloat16* __restrict__ w, uint8_t* __restrict__ dst) {
  const int tile = blockIdx.x, par = tile >> 1, grp = tile & 1;
  for (int i = threadIdx.x; i < 128 * 8; i += blockDim.x) {
    const int row = i >> 3, j = i & 7;           // 16-byte chunk j = channels 8j .. 8j+7 of the 64-wide K
    const int blk = row >> 6, co = row & 63;
    const int ky = par == 0 ? (blk == 0 ? 2 : 0) : (blk == 0 ? 3 : 1);
    const int half = j >> 2;                      // which 32-channel half of the pair row this chunk multiplies
    const int kx = grp == 0 ? (half == 0 ? 1 : 2) : (half == 0 ? 3 : 0);
    const int ci = (j & 3) * 8;
    const uint4 v = *reinterpret_cast<const uint4*>(w + (int64_t)co * 16 * kDwCin + (ky * 4 + kx) * kDwCin + ci);
    *reinterpret_cast<uint4*>(dst + (int64_t)tile * kDwWTile + row * 128 + ((j ^ (row & 7)) << 4)) = v;
  }
}

struct DwUnit {
  int b, w0, t0, rows;   // output coordinates
  __device__ __forceinline__ void set(const DwParams& p, int u) {
    const int per_b = ((p.W / 2) >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    w0 = wb << 7;
    rows = min(p.RT, p.T / 2 - t0);
  }
};

__global__ void __launch_bounds__(kDwThreads, 1)
conv_down_wide_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ DwParams p) {
  constexpr int NS = kDwSlots, NB = kDwNB, CO = kDwCout;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t w_base = smem_base;
  const uint32_t a_base = smem_base + kDwWBytes;
  const uint32_t aux = a_base + NS * kDwSlotBytes;
  uint8_t* aux_gen = smem_gen + kDwWBytes + NS * kDwSlotBytes;
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

  if (warp == 0) {
    // ================================================================ TMA producer: input rows 2 t0 - 1 .. 2 (t0 + rows) of every unit
    if (elect_one()) {
      mbar_arrive_expect_tx(bar_w, (uint32_t)kDwWBytes);
      bulk_load_g2s(w_base, p.wimg, (uint32_t)kDwWBytes, bar_w);
      uint32_t cnt = 0;
      DwUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < 2 * u.rows + 2; ++rr, ++cnt) {
          const uint32_t slot = cnt % NS, use = cnt / NS;
          mbar_wait(bar_empty + 8 * slot, (use & 1) ^ 1, 51);
          mbar_arrive_expect_tx(bar_full + 8 * slot, (uint32_t)(kWidePos * 128));
          tma_load_4d(a_base + slot * kDwSlotBytes, &tmA, bar_full + 8 * slot, 0, u.w0 - 1, 2 * u.t0 - 1 + rr, u.b);
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread)
    if (elect_one()) {
      mbar_wait(bar_w, 0, 52);
      tc_fence_after();
      const uint64_t desc_hi = make_kmajor_desc(0, 1024, 2);             // SWIZZLE_128B, 8-row group pitch 1024 B
      const uint32_t w_lo = (w_base & 0x3FFFFu) >> 4;
      constexpr uint32_t kBlk16 = (64 * 128) >> 4;                       // second 64-row block of a weight tile
      uint32_t cnt = 0, otile = 0;
      DwUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < 2 * u.rows + 2; ++rr) {
          const uint32_t slot = (cnt + rr) % NS;
          const int i = rr >> 1, par = rr & 1;                           // local output blocks (i - 1, i); par 0: ky (2, 0), par 1: ky (3, 1)
          const bool vlo = i >= 1, vhi = i < u.rows;
          const uint32_t blo = (otile + i + NB - 1) % NB, bhi = (otile + i) % NB;
          const bool first_hi = vhi && par == 0;                         // block i receives its first contribution from this row
          const bool merged = vlo && vhi && bhi != 0;                    // adjacent in the TMEM ring
          mbar_wait(bar_full + 8 * slot, ((cnt + rr) / NS) & 1, 53);
          if (first_hi) mbar_wait(bar_aempty + 8 * bhi, (((otile + i) / NB) & 1) ^ 1, 54);
          tc_fence_after();
          const uint32_t a_lo = ((a_base + slot * kDwSlotBytes) & 0x3FFFFu) >> 4;
          const uint32_t wt = w_lo + (uint32_t)(par * 2) * (kDwWTile >> 4);
          const uint32_t t_lo = tmem_base + blo * CO, t_hi = tmem_base + bhi * CO;
          const uint32_t id1 = make_idesc_bf16(kTileM, (uint32_t)CO), id2 = make_idesc_bf16(kTileM, (uint32_t)(2 * CO));
          // k-steps: centre group (pair f', A shift 1 row = 8 units), k = 0..3; kx = 3 (pair f'+1, shift 2, bytes 0..63), k = 0,1;
          //          kx = 0 (pair f'-1, shift 0, bytes 64..127), k = 2,3.  Weight tile 0 = centre, tile 1 = sides.
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            const int grp = s < 4 ? 0 : 1;
            const int k = s < 4 ? s : (s - 4);
            const int shift = s < 4 ? 1 : (s < 6 ? 2 : 0);
            const uint64_t ad = desc_hi | (uint64_t)(a_lo + (uint32_t)(shift * 8 + 2 * k));
            const uint32_t bl = wt + (uint32_t)grp * (kDwWTile >> 4) + 2 * k;
            if (s == 0 && first_hi) {
              if (vlo) umma_bf16(t_lo, ad, desc_hi | (uint64_t)bl, id1, 1u);
              umma_bf16(t_hi, ad, desc_hi | (uint64_t)(bl + kBlk16), id1, 0u);
            } else if (merged) {
              umma_bf16(t_lo, ad, desc_hi | (uint64_t)bl, id2, 1u);
            } else {
              if (vlo) umma_bf16(t_lo, ad, desc_hi | (uint64_t)bl, id1, 1u);
              if (vhi) umma_bf16(t_hi, ad, desc_hi | (uint64_t)(bl + kBlk16), id1, 1u);
            }
          }
          umma_commit(bar_empty + 8 * slot);
          if (par == 1 && vlo) umma_commit(bar_afull + 8 * blo);          // output row i - 1 has all four kernel rows
        }
        cnt += 2 * u.rows + 2;
        otile += u.rows;
      }
    }
  } else {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4)
    const int q = warp & 3;
    const int m = q * 32 + lane;
    float* wb = bias_smem + q * 256;
    uint8_t* stage = aux_gen + 1024 + 4096 + q * 2048;
    ConvEpiState<CO> st;
    st.reset();
    st.cur_batch = -1;
    const int To = p.T / 2, Wo = p.W / 2;
    uint32_t tile = 0;
    DwUnit u;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      u.set(p, unit);
      if (u.b != st.cur_batch) {
        st.flush(p.stats, lane);
        st.cur_batch = u.b;
        conv_epilogue_bias<CO>(wb, p.bias, nullptr, u.b, 0, 0, lane);
      }
      for (int ot = 0; ot < u.rows; ++ot, ++tile) {
        const uint32_t buf = tile % NB, use = tile / NB;
        const int64_t addr0 = ((((int64_t)u.b * To + (u.t0 + ot)) * Wo) + u.w0 + m) * CO;
        mbar_wait(bar_afull + 8 * buf, use & 1, 55);
        tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * CO;
        conv_epilogue_tile<CO, CO>(trow, wb, 0, nullptr, p.out, addr0, true, p.stats != nullptr, st, stage);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
      }
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
