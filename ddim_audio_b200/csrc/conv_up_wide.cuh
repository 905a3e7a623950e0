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
// output rows): the skip tile of an output row ([32 pixel pairs][128 B] per warp) arrives by TMA into one of the warp's two
// staging tiles a tile ahead, bias + skip are added in place, the tile leaves by ONE TMA store (r02: the per-thread staged
// loads / stores of the skip and output rows kept this kernel at 0.38 of HBM).
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kUwThreads = 320;
constexpr int kUwCin = 64, kUwCout = 32;
constexpr int kUwSlotBytes = ((kWidePos * 128 + 1023) / 1024) * 1024;   // 17408
constexpr int kUwSlots = 3;
constexpr int kUwNB = 8;                               // output rows in flight in TMEM (64 columns each)
constexpr int kUwWTile = 256 * 128;                    // [4 ky][2 s][32 co] rows x 64 ci
constexpr int kUwWBytes = 3 * kUwWTile;
constexpr int kUwStage = 32 * 128;                     // one epilogue warp's tile: 32 pixel pairs x [s][co] (128 B), SWIZZLE_128B image
constexpr int kUwAux = 1024 + 4096 + 8 * 2 * kUwStage; // barriers | 8 bias copies | 8 warps x 2 staging tiles
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
conv_up_wide_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmS,
                    const __grid_constant__ UwParams p) {
  constexpr int NS = kUwSlots, NB = kUwNB, NA = 2 * kUwCout;   // accumulator block width
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t w_base = smem_base;
  const uint32_t a_base = smem_base + kUwWBytes;
  const uint32_t aux = a_base + NS * kUwSlotBytes;
  uint8_t* aux_gen = smem_gen + kUwWBytes + NS * kUwSlotBytes;
  // aux: full[8] @0, empty[8] @64, acc_full[8] @128, acc_empty[8] @192, wfull @256, tmem ptr @264, skip[8 warps][2] @320
  const uint32_t bar_full = aux, bar_empty = aux + 64, bar_afull = aux + 128, bar_aempty = aux + 192, bar_w = aux + 256, bar_skip = aux + 320;
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
    for (int i = 0; i < 16; ++i) mbar_init(bar_skip + 8 * i, 1);
    tma_prefetch_desc(&tmO);
    tma_prefetch_desc(&tmS);
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
    float* wb = bias_smem + (warp - 2) * 128;
    uint8_t* stage_gen = aux_gen + 1024 + 4096 + (warp - 2) * 2 * kUwStage;
    const uint32_t stage = smem_u32(stage_gen);
    const uint32_t bsk = bar_skip + 16 * (warp - 2);
    const uint32_t phase = (uint32_t)(lane & 7);
    const bool has_skip = p.skip != nullptr, want_stats = p.stats != nullptr;
    ConvEpiState<kUwCout> st;
    st.reset();
    st.cur_batch = -1;
    if (grp == 0 && lane == 0) mbar_arrive(bar_aempty + 0);   // same for the "accumulator free" barrier of ring position 0 (4 arrivals)

    // this warp's tiles in order: (unit, local output row lo) with lo of this warp's parity group
    struct Walk {
      int unit, lo, nout;
      uint32_t g0;
      UwUnit u;
      __device__ __forceinline__ void begin(const UwParams& p, uint32_t grp) {
        unit = blockIdx.x; g0 = 0;
        enter(p, grp);
      }
      __device__ __forceinline__ void enter(const UwParams& p, uint32_t grp) {
        while (unit < p.num_units) {
          u.set(p, unit);
          nout = 2 * u.rows;
          lo = (int)((grp - g0) & 1u);
          if (lo < nout) return;
          g0 += (uint32_t)nout;
          unit += gridDim.x;
        }
      }
      __device__ __forceinline__ bool valid(const UwParams& p) const { return unit < p.num_units; }
      __device__ __forceinline__ void next(const UwParams& p, uint32_t grp) {
        lo += 2;
        if (lo >= nout) {
          g0 += (uint32_t)nout;
          unit += gridDim.x;
          enter(p, grp);
        }
      }
    };
    Walk cur, nxt;
    cur.begin(p, grp);
    nxt = cur;
    if (nxt.valid(p)) nxt.next(p, grp);
    uint32_t it = 0;
    auto load_skip = [&](const Walk& w, uint32_t b) {
      if (has_skip && elect_one()) {
        mbar_arrive_expect_tx(bsk + 8 * b, (uint32_t)kUwStage);
        tma_load_4d(stage + b * kUwStage, &tmS, bsk + 8 * b, 0, w.u.w0 + q * 32, 2 * w.u.t0 + w.lo, w.u.b);
      }
    };
    if (cur.valid(p)) load_skip(cur, 0);
    for (; cur.valid(p); ++it) {
      const uint32_t sb = it & 1u;
      if (cur.u.b != st.cur_batch) {
        st.flush(p.stats, lane);
        st.cur_batch = cur.u.b;
        conv_epilogue_bias<kUwCout>(wb, p.bias, nullptr, cur.u.b, 0, 0, lane);
      }
      // the other staging tile was the source of the previous tile's store: once that has been read, the next tile's skip may land in it
      if (it > 0) {
        bulk_wait_group_read<0>();
        __syncwarp();
      }
      if (nxt.valid(p)) load_skip(nxt, sb ^ 1u);
      const uint32_t g = cur.g0 + (uint32_t)cur.lo + 1;
      const uint32_t buf = g % NB, use = g / NB;
      mbar_wait(bar_afull + 8 * buf, use & 1, 65);
      tc_fence_after();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * NA;
      uint8_t* srow = stage_gen + sb * kUwStage + lane * 128;
#pragma unroll
      for (int c0 = 0; c0 < NA; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(trow + (uint32_t)c0, v);
        tmem_ld_wait();
        if (c0 + 32 == NA) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
        }
        uint64_t f[16];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(wb + ((c0 + 2 * j) % kUwCout));
          f[j] = fadd2(pack2u(v[2 * j], v[2 * j + 1]), b4.x);
          f[j + 1] = fadd2(pack2u(v[2 * j + 2], v[2 * j + 3]), b4.y);
        }
        if (has_skip) {
          if (c0 == 0) {
            mbar_wait(bsk + 8 * sb, (it >> 1) & 1u, 66);   // staging tile sb is used by every second tile
          }
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) {
            const uint4 a = *reinterpret_cast<const uint4*>(srow + ((((uint32_t)(c0 >> 3) + qq) ^ phase) << 4));
            f[4 * qq] = fadd2(f[4 * qq], bf16x2_to_f32x2(a.x));
            f[4 * qq + 1] = fadd2(f[4 * qq + 1], bf16x2_to_f32x2(a.y));
            f[4 * qq + 2] = fadd2(f[4 * qq + 2], bf16x2_to_f32x2(a.z));
            f[4 * qq + 3] = fadd2(f[4 * qq + 3], bf16x2_to_f32x2(a.w));
          }
        }
#pragma unroll
        for (int qq = 0; qq < 4; ++qq)
          *reinterpret_cast<uint4*>(srow + ((((uint32_t)(c0 >> 3) + qq) ^ phase) << 4)) =
              make_uint4(f32x2_to_bf16x2(f[4 * qq]), f32x2_to_bf16x2(f[4 * qq + 1]), f32x2_to_bf16x2(f[4 * qq + 2]), f32x2_to_bf16x2(f[4 * qq + 3]));
        if (want_stats) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int gi = ((c0 + 2 * j) % kUwCout) / (kUwCout / 8);   // compile-time after unrolling
            st.s1[gi] = fadd2(st.s1[gi], f[j]);
            st.s2[gi] = ffma2(f[j], f[j], st.s2[gi]);
          }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      tma_store_4d_e(&tmO, stage + sb * kUwStage, 0, cur.u.w0 + q * 32, 2 * cur.u.t0 + cur.lo, cur.u.b);
      bulk_commit_group();
      cur = nxt;
      if (nxt.valid(p)) nxt.next(p, grp);
    }
    st.flush(p.stats, lane);
    bulk_wait_group<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace ddimb
