// Weight gradient of the 3x3 convolutions of the two widest levels (C = 32, 64; W a multiple of 128): autograd of reference
// models/diffusion.py:28-40,  dw[co][ci][ky][kx] += sum_{b,t,w} dy[b,t,w,co] * x[b,t+ky-1,w+kx-1,ci].
//
// The contraction runs over 10^8 positions against a 9 x C x C result: the op is HBM-bound on reading x and dy ONCE (arithmetic
// intensity 144 / 288 flop/B).  The generic split-K kernel (gemm_tn.cuh) re-reads both tensors for each of the 9 taps.  Here a CTA
// walks a 128-column strip down consecutive rows like the forward kernel (conv_wide.cuh): every x row strip (130 positions: halo
// columns, TMA zero fill outside the image) and every dy row strip (128 positions) is read once into a ring slot, and the
// accumulators of ALL NINE taps stay in tensor memory for the CTA's whole lifetime; they are added to dw once at the end.
//   * both operands are MN-major (channels contiguous per position, exactly as TMA delivers NHWC rows); K = the 128 positions of
//     the dy strip, 16 per MMA.
//   * kx lives in the M dimension: the A descriptor's leading-dimension offset is ONE POSITION ROW, so its 128 / C channel atoms are
//     the x strip shifted by 0, 1, ... positions: D row (kx, ci) = sum_p x[p + kx][ci] dy[p][.].  C = 32: one MMA covers kx = 0..2
//     (the fourth atom is a phantom whose D rows are never read); C = 64: two MMAs (kx = 0, 1 and kx = 2 + phantom).
//   * ky lives in the N dimension: x row r meets dy rows r-2, r-1, r (ky = 2, 1, 0), which sit in CONSECUTIVE ring slots, so the B
//     descriptor's leading-dimension offset is the slot pitch and one MMA of N = 3 C covers all three (split in two where the ring
//     wraps, trimmed at the first / last rows of a unit).
//   so an x row costs 8 (C = 32) / 16 (C = 64) MMAs for all nine taps.  (The first version put kx in N through position-shifted dy
//   atoms and ran ky as separate MMAs: 27 MMAs of N = 96 per row at ~90 cycles each, issue-bound at 0.33 of HBM: r02 experiments 8.)
//   * the accumulators start zeroed (tcgen05.st), every MMA accumulates.
// Warp roles (256 threads): 0 = TMA producer, 1 = MMA issuer, 4-7 = accumulator zeroing and the final reduction (TMEM lane quarters).
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kWgThreads = 256;
constexpr int kWgXPos = 130;                           // x positions per slot: w0 - 1 .. w0 + 128

template <int C>
struct WgCfg {
  static constexpr int kRowBytes = 2 * C;
  static constexpr int kAtoms = 128 / C;               // position-shifted channel atoms per A operand (4 / 2)
  static constexpr int kPasses = (3 + kAtoms - 1) / kAtoms;   // MMAs per K step (1 / 2)
  static constexpr int kXSlot = ((kWgXPos + 6) * kRowBytes + 1023) / 1024 * 1024;   // + the phantom atoms' overrun (C = 32: position 130; C = 64: 131)
  static constexpr int kYSlot = 128 * kRowBytes;
  static constexpr int kXSlots = (C == 32) ? 10 : 6, kYSlots = (C == 32) ? 10 : 6;
  static constexpr int kTmemCols = (C == 32) ? 128 : 512;     // kPasses x 3 C accumulator columns (96 / 384), rounded up
  static constexpr int kSwizzle = (C == 64) ? 2 : 4;
  static constexpr int kSmem = 1024 + kXSlots * kXSlot + kYSlots * kYSlot + 1024;
};

struct WgParams {
  int32_t B, T, W;
  int32_t RT, units_t, num_units;
  float* dw;              // fp32 [C][C][3][3], accumulated
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

__device__ __forceinline__ uint64_t wg_desc_hi(uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t swizzle_code) {
  return ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)swizzle_code << 61);
}

template <int C>
__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_wide_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmY, const __grid_constant__ WgParams p) {
  using Cfg = WgCfg<C>;
  constexpr int NX = Cfg::kXSlots, NY = Cfg::kYSlots, RB = Cfg::kRowBytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t x_base = smem_base, y_base = smem_base + NX * Cfg::kXSlot;
  const uint32_t aux = y_base + NY * Cfg::kYSlot;
  uint8_t* aux_gen = smem_gen + NX * Cfg::kXSlot + NY * Cfg::kYSlot;
  // aux: x_full[16] @0, x_empty[16] @128, y_full[16] @256, y_empty[16] @384, done @512, zeroed @520, tmem ptr @528
  const uint32_t bar_xf = aux, bar_xe = aux + 128, bar_yf = aux + 256, bar_ye = aux + 384, bar_done = aux + 512, bar_zero = aux + 520;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 528);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // the x slots' tail rows (beyond the 130 TMA writes, read only by the phantom atoms) must hold finite numbers: NaN bit patterns in
  // fresh shared memory would stay confined to D rows nobody reads, but zero them anyway
  for (int i = threadIdx.x; i < NX * Cfg::kXSlot / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem_gen)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmY);
    for (int s = 0; s < NX; ++s) { mbar_init(bar_xf + 8 * s, 1); mbar_init(bar_xe + 8 * s, 1); }
    for (int s = 0; s < NY; ++s) { mbar_init(bar_yf + 8 * s, 1); mbar_init(bar_ye + 8 * s, 1); }
    mbar_init(bar_done, 1);
    mbar_init(bar_zero, 4);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 528, (uint32_t)Cfg::kTmemCols);
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
            mbar_arrive_expect_tx(bar_xf + 8 * slot, (uint32_t)(kWgXPos * RB));
            tma_load_4d(x_base + slot * Cfg::kXSlot, &tmX, bar_xf + 8 * slot, 0, u.w0 - 1, u.t0 - 1 + rr, u.b);
            ++xc;
          }
          if (rr < u.rows) {
            const uint32_t slot = yc % NY, use = yc / NY;
            mbar_wait(bar_ye + 8 * slot, (use & 1) ^ 1, 82);
            mbar_arrive_expect_tx(bar_yf + 8 * slot, (uint32_t)Cfg::kYSlot);
            tma_load_4d(y_base + slot * Cfg::kYSlot, &tmY, bar_yf + 8 * slot, 0, u.w0, u.t0 + rr, u.b);
            ++yc;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread)
    if (elect_one()) {
      // MN-major operands, atom = [8 positions][C channels], SBO = 8 position rows.  A: atoms one position row apart (kx);
      // B: atoms one ring slot apart (ky)
      const uint64_t hi_a = wg_desc_hi((uint32_t)RB, 8u * RB, (uint32_t)Cfg::kSwizzle);
      const uint64_t hi_b = wg_desc_hi((uint32_t)Cfg::kYSlot, 8u * RB, (uint32_t)Cfg::kSwizzle);
      mbar_wait(bar_zero, 0, 86);
      tc_fence_after();
      uint32_t xc = 0, yc = 0;
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
            // column block j <-> dy row rr - 2 + j <-> ky = 2 - j; the valid blocks jf..jl are contiguous
            const int jf = rr >= 2 ? 0 : 2 - rr, jl = rr < u.rows ? 2 : u.rows + 1 - rr;
            if (jf <= jl) {
              const uint32_t s0 = (yc + (uint32_t)(rr - 2 + jf)) % NY;      // ring slot of block jf
              const int n1 = min(jl - jf + 1, NY - (int)s0);                // blocks before the ring wraps
              const int n2 = jl - jf + 1 - n1;
              const uint32_t a_lo = ((x_base + xs * Cfg::kXSlot) & 0x3FFFFu) >> 4;
              const uint32_t b1_lo = ((y_base + s0 * Cfg::kYSlot) & 0x3FFFFu) >> 4, b2_lo = (y_base & 0x3FFFFu) >> 4;
              const uint32_t id1 = make_idesc_bf16(128, (uint32_t)(n1 * C)) | (1u << 15) | (1u << 16);
              const uint32_t id2 = make_idesc_bf16(128, (uint32_t)((n2 > 0 ? n2 : 1) * C)) | (1u << 15) | (1u << 16);
#pragma unroll
              for (int ps = 0; ps < Cfg::kPasses; ++ps) {
                const uint32_t acc = tmem_base + (uint32_t)(ps * 3 * C + jf * C);
                const uint32_t a_ps = a_lo + (uint32_t)((ps * Cfg::kAtoms * RB) >> 4);   // pass ps starts kAtoms positions further on
#pragma unroll
                for (int s = 0; s < 8; ++s) {
                  const uint32_t koff = (uint32_t)(s * 16 * RB) >> 4;
                  umma_bf16(acc, hi_a | (uint64_t)(a_ps + koff), hi_b | (uint64_t)(b1_lo + koff), id1, 1u);
                  if (n2 > 0) umma_bf16(acc + (uint32_t)(n1 * C), hi_a | (uint64_t)(a_ps + koff), hi_b | (uint64_t)(b2_lo + koff), id2, 1u);
                }
              }
            }
          }
          umma_commit(bar_xe + 8 * xs);                      // x row consumed
          if (rr >= 2) umma_commit(bar_ye + 8 * ((yc + (uint32_t)(rr - 2)) % NY));   // dy row rr - 2 has met its three x rows
        }
        xc += (uint32_t)(u.rows + 2);
        yc += (uint32_t)u.rows;
      }
      umma_commit(bar_done);
    }
  } else if (warp >= 4) {
    // ================================================================ zero the accumulators, wait, add them to dw
    const int q = warp & 3;
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll
    for (int c0 = 0; c0 < Cfg::kPasses * 3 * C; c0 += 32) tmem_st_zero32(trow + (uint32_t)c0);
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_zero);
    mbar_wait(bar_done, 0, 85);
    tc_fence_after();
    // TMEM lane (q, lane) of pass ps = D row (kx, ci): kx = ps * kAtoms + (q * 32 + lane) / C, ci = (q * 32 + lane) % C.
    // The accumulators are laid out in shared memory as dw is ([co][ci][ky][kx], 9 C^2 floats over the idle x / dy rings: every MMA has
    // completed) and then added to dw with consecutive lanes on consecutive addresses: 4 sectors per warp atomic instead of the 32 a
    // lane-per-ci pattern touches (148 CTAs x 9 C^2 scattered atomics were a fixed 40-60 us of every call).
    float* stage = reinterpret_cast<float*>(smem_gen);
    static_assert(9 * C * C * 4 <= NX * Cfg::kXSlot + NY * Cfg::kYSlot, "staging tile fits in the rings");
    const int row = q * 32 + lane;
    const int ci = row % C;
#pragma unroll 1
    for (int ps = 0; ps < Cfg::kPasses; ++ps) {
      const int kx = ps * Cfg::kAtoms + row / C;
      if (kx > 2) continue;                                  // phantom atoms (warp-uniform: C >= 32)
#pragma unroll 1
      for (int c0 = 0; c0 < 3 * C; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(trow + (uint32_t)(ps * 3 * C + c0), v);
        tmem_ld_wait();
        const int ky = 2 - c0 / C, co0 = c0 % C;
#pragma unroll
        for (int i = 0; i < 16; ++i) stage[((co0 + i) * C + ci) * 9 + ky * 3 + kx] = __uint_as_float(v[i]);
      }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");          // the four reduction warps
    for (int i = (warp - 4) * 32 + lane; i < 9 * C * C; i += 128) atomicAdd(p.dw + i, stage[i]);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)Cfg::kTmemCols);
  }
}

}  // namespace ddimb
