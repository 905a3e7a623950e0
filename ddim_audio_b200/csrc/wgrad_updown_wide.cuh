// Weight gradients of the two widest strided convolutions (level 0 <-> 1): Downsample Conv2d(32 -> 64, k4 s2 p1) and Upsample
// ConvTranspose2d(64 -> 32, k4 s2 p1), autograd of reference models/diffusion.py:59-78.  Both pair a high-resolution 32-channel tensor
// H [B][T][W][32] with a low-resolution 64-channel tensor L [B][T/2][W/2][64] and accumulate
//     G[c64][c32][ky][kx] += sum_{b,t',f'} L[b,t',f'][c64] * H[b, 2t'+ky-1, 2f'+kx-1][c32]
// (Downsample: H = x, L = dy, G = dw[co][ci]; Upsample: H = dy, L = x, G = dw[ci][co]).  The generic split-K kernel runs the 16 taps
// as 16 jobs that each re-read H and L (0.15 of HBM).  Here, as in wgrad_wide.cuh, a CTA walks a strip of 128 low-resolution columns
// down the rows, reads every H and L row strip once through TMA rings and keeps the accumulators of ALL SIXTEEN taps in tensor memory:
//   * H rows are viewed space-to-depth: pixel pair q = (2q, 2q+1) is one 128-byte row of 64 channels; both operands are MN-major.
//   * kx lives in M: the A descriptor's leading-dimension offset is one pair row, so its two 64-channel atoms are pairs q and q + 1.
//     Started at pair f'-1 they give kx = 0 (odd half of f'-1), 1, 2 (both halves of f'); started at f'+1, kx = 3 (even half) -- two
//     MMAs per K step, 5 of their 8 row blocks of 32 used.
//   * ky lives in N and in the row parity: H row r meets the two L rows t' with 2t'+ky-1 = r -- consecutive ring slots, so the B
//     descriptor's leading-dimension offset is the slot pitch (N = 2 x 64); odd r: ky = (2, 0), even r: ky = (3, 1), each parity with
//     its own accumulators.  4 accumulator tiles of 128 columns = all 512 TMEM columns.
//   so an H row costs 16 MMAs of N = 128 for all sixteen taps.  The accumulators start zeroed; at the end they are laid out in shared
//   memory as G is and added with coalesced atomics.
// Warp roles (256 threads): 0 = TMA producer, 1 = MMA issuer, 4-7 = accumulator zeroing and the final reduction.
#pragma once
#include "wgrad_wide.cuh"

namespace ddimb {

constexpr int kUdThreads = 256;
constexpr int kUdHPairs = 130;                          // pairs per H slot: f'0 - 1 .. f'0 + 128
constexpr int kUdHSlot = 17408;                         // 130 x 128 B + the phantom atom's overrun, rounded to 1024
constexpr int kUdLSlot = 128 * 128;
constexpr int kUdHSlots = 6, kUdLSlots = 6;
constexpr int kUdSmem = 1024 + kUdHSlots * kUdHSlot + kUdLSlots * kUdLSlot + 1024;

struct UdParams {
  int32_t B, T, W;        // extents of H (T, W even; W / 2 a multiple of 128)
  int32_t RT, units_t, num_units;   // units of RT low-resolution rows
  float* g;               // fp32 [64][32][4][4], accumulated
};

struct UdUnit {
  int b, f0, t0, rows;    // low-resolution coordinates
  __device__ __forceinline__ void set(const UdParams& p, int u) {
    const int per_b = ((p.W / 2) >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    f0 = wb << 7;
    rows = min(p.RT, p.T / 2 - t0);
  }
};

__global__ void __launch_bounds__(kUdThreads, 1)
wgrad_updown_wide_kernel(const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmL, const __grid_constant__ UdParams p) {
  constexpr int NH = kUdHSlots, NL = kUdLSlots;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t h_base = smem_base, l_base = smem_base + NH * kUdHSlot;
  const uint32_t aux = l_base + NL * kUdLSlot;
  uint8_t* aux_gen = smem_gen + NH * kUdHSlot + NL * kUdLSlot;
  // aux: h_full[8] @0, h_empty[8] @64, l_full[8] @128, l_empty[8] @192, done @256, zeroed @264, tmem ptr @272
  const uint32_t bar_hf = aux, bar_he = aux + 64, bar_lf = aux + 128, bar_le = aux + 192, bar_done = aux + 256, bar_zero = aux + 264;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 272);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // rows beyond the 130 TMA writes of an H slot are read by the phantom atom only; keep them finite
  for (int i = threadIdx.x; i < NH * kUdHSlot / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmH);
    tma_prefetch_desc(&tmL);
    for (int s = 0; s < NH; ++s) { mbar_init(bar_hf + 8 * s, 1); mbar_init(bar_he + 8 * s, 1); }
    for (int s = 0; s < NL; ++s) { mbar_init(bar_lf + 8 * s, 1); mbar_init(bar_le + 8 * s, 1); }
    mbar_init(bar_done, 1);
    mbar_init(bar_zero, 4);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 272, 512u);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ TMA producer: per unit H rows 2 t0 - 1 .. 2 (t0 + rows), L rows t0 .. t0 + rows - 1
    if (elect_one()) {
      uint32_t hc = 0, lc = 0;
      UdUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < 2 * u.rows + 2; ++rr) {
          if ((rr & 1) == 0 && (rr >> 1) < u.rows) {           // L row rr / 2 is first needed by H row rr
            const uint32_t slot = lc % NL, use = lc / NL;
            mbar_wait(bar_le + 8 * slot, (use & 1) ^ 1, 92);
            mbar_arrive_expect_tx(bar_lf + 8 * slot, (uint32_t)kUdLSlot);
            tma_load_4d(l_base + slot * kUdLSlot, &tmL, bar_lf + 8 * slot, 0, u.f0, u.t0 + (rr >> 1), u.b);
            ++lc;
          }
          const uint32_t slot = hc % NH, use = hc / NH;
          mbar_wait(bar_he + 8 * slot, (use & 1) ^ 1, 91);
          mbar_arrive_expect_tx(bar_hf + 8 * slot, (uint32_t)(kUdHPairs * 128));
          tma_load_4d(h_base + slot * kUdHSlot, &tmH, bar_hf + 8 * slot, 0, u.f0 - 1, 2 * u.t0 - 1 + rr, u.b);
          ++hc;
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread)
    if (elect_one()) {
      // MN-major SWIZZLE_128B operands, atom = [8 positions][64 channels], SBO = 8 rows.  A: atoms one pair row apart; B: one ring slot apart
      const uint64_t hi_a = wg_desc_hi(128u, 1024u, 2u);
      const uint64_t hi_b = wg_desc_hi((uint32_t)kUdLSlot, 1024u, 2u);
      mbar_wait(bar_zero, 0, 96);
      tc_fence_after();
      uint32_t hc = 0, lc = 0;
      UdUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < 2 * u.rows + 2; ++rr) {
          const uint32_t hs = (hc + rr) % NH;
          mbar_wait(bar_hf + 8 * hs, ((hc + rr) / NH) & 1, 93);
          if ((rr & 1) == 0 && (rr >> 1) < u.rows) mbar_wait(bar_lf + 8 * ((lc + (rr >> 1)) % NL), ((lc + (rr >> 1)) / NL) & 1, 94);
          tc_fence_after();
          const int r = 2 * u.t0 - 1 + rr;                        // H row
          // column block j <-> L row i_lo + j of the unit, i_lo = rr / 2 - 1; ky = (2, 0) for even rr (odd r), (3, 1) for odd rr
          const int i_lo = (rr >> 1) - 1;
          const int jf = i_lo < 0 ? 1 : 0, jl = (i_lo + 1 < u.rows) ? 1 : 0;
          if (r >= 0 && r < p.T && jf <= jl) {
            const uint32_t s0 = (lc + (uint32_t)(i_lo + jf)) % NL;
            const int n1 = min(jl - jf + 1, NL - (int)s0), n2 = jl - jf + 1 - n1;
            const uint32_t a_lo = ((h_base + hs * kUdHSlot) & 0x3FFFFu) >> 4;
            const uint32_t b1_lo = ((l_base + s0 * kUdLSlot) & 0x3FFFFu) >> 4, b2_lo = (l_base & 0x3FFFFu) >> 4;
            const uint32_t id1 = make_idesc_bf16(128, (uint32_t)(n1 * 64)) | (1u << 15) | (1u << 16);
            const uint32_t id2 = make_idesc_bf16(128, 64u) | (1u << 15) | (1u << 16);
#pragma unroll
            for (int ps = 0; ps < 2; ++ps) {
              const uint32_t acc = tmem_base + (uint32_t)(((rr & 1) * 2 + ps) * 128 + jf * 64);
              const uint32_t a_ps = a_lo + (uint32_t)((ps * 2 * 128) >> 4);        // pass 1 starts two pairs further on
#pragma unroll
              for (int s = 0; s < 8; ++s) {
                const uint32_t koff = (uint32_t)(s * 16 * 128) >> 4;
                umma_bf16(acc, hi_a | (uint64_t)(a_ps + koff), hi_b | (uint64_t)(b1_lo + koff), id1, 1u);
                if (n2 > 0) umma_bf16(acc + (uint32_t)(n1 * 64), hi_a | (uint64_t)(a_ps + koff), hi_b | (uint64_t)(b2_lo + koff), id2, 1u);
              }
            }
          }
          umma_commit(bar_he + 8 * hs);                           // H row consumed
          if ((rr & 1) == 1 && rr >= 3) umma_commit(bar_le + 8 * ((lc + (uint32_t)((rr - 3) >> 1)) % NL));   // L row (rr - 3) / 2 has met its four H rows
        }
        hc += (uint32_t)(2 * u.rows + 2);
        lc += (uint32_t)u.rows;
      }
      umma_commit(bar_done);
    }
  } else if (warp >= 4) {
    // ================================================================ zero the accumulators, wait, add them to G
    const int q = warp & 3;
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll
    for (int c0 = 0; c0 < 512; c0 += 32) tmem_st_zero32(trow + (uint32_t)c0);
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_zero);
    mbar_wait(bar_done, 0, 95);
    tc_fence_after();
    // TMEM lane quarter q of pass ps: pass 0: q = 1, 2, 3 <-> kx = 0, 1, 2; pass 1: q = 0 <-> kx = 3 (lane = c32); the other quarters are
    // unused halves / the phantom atom.  Column c of parity pr: j = c / 64 (L row), c64 = c % 64; ky = pr ? (j ? 1 : 3) : (j ? 0 : 2).
    float* stage = reinterpret_cast<float*>(smem_gen);            // [c64][c32][ky][kx] over the idle rings
    static_assert(64 * 32 * 16 * 4 <= NH * kUdHSlot + NL * kUdLSlot, "staging tile fits in the rings");
#pragma unroll 1
    for (int ps = 0; ps < 2; ++ps) {
      const int kx = ps == 0 ? q - 1 : (q == 0 ? 3 : -1);
      if (kx < 0) continue;
#pragma unroll 1
      for (int pr = 0; pr < 2; ++pr) {
#pragma unroll 1
        for (int c0 = 0; c0 < 128; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(trow + (uint32_t)((pr * 2 + ps) * 128 + c0), v);
          tmem_ld_wait();
          const int j = c0 >> 6, ky = pr ? (j ? 1 : 3) : (j ? 0 : 2);
#pragma unroll
          for (int i = 0; i < 16; ++i) stage[(((c0 & 63) + i) * 32 + lane) * 16 + ky * 4 + kx] = __uint_as_float(v[i]);
        }
      }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");                // the four reduction warps
    for (int i = (warp - 4) * 32 + lane; i < 64 * 32 * 16; i += 128) atomicAdd(p.g + i, stage[i]);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512u);
  }
}

}  // namespace ddimb
