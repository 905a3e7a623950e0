// 3x3 convolution of the two widest U-Net levels (C = 32 / 64 channels, W a multiple of 128), where the work is HBM-bound:
// reference models/diffusion.py:42-56 -- GroupNorm (+SiLU) -> Conv2d 3x3 -> (+temb | +bias) -> SiLU, in ONE pass over x.
//
// A CTA walks a strip of 128 output columns down RT consecutive rows.  Every input row strip (130 positions x C channels,
// halo columns included, TMA zero fill outside the image) is read from HBM/L2 ONCE into a ring slot; the GroupNorm affine
// (+SiLU) of the block's first half is applied in place by four transform warps (halo positions forced back to zero, which
// is what the reference's zero padding of the ACTIVATED tensor means); then all 9 taps are tcgen05.mma's whose A descriptors
// simply start kx rows further into slot (t + ky - 1) -- the swizzle is a function of the absolute shared-memory address,
// so a row-shifted start needs no re-layout (tools/umma_probe.cu).  The 9 weight tiles stay resident in shared memory.
// Warp roles (512 threads, 128 registers each): 0 = TMA producer, 1 = MMA issuer, 2-9 = epilogue (two warps per TMEM lane quarter, alternating
// output rows), 10-15 = transform (every sixth row strip each).  The epilogue and transform warps are latency-bound
// dependent chains (ncu: one warp of each per scheduler issued every ~6 cycles), so each scheduler gets two of each.
#pragma once
#include "umma_gemm.cuh"

namespace ddimb {

#ifndef DDIMB_WIDE_BULKST
#define DDIMB_WIDE_BULKST 1
#endif
// Accumulator protocol: the epilogue warps hand every accumulator block back ZEROED (tcgen05.st after their tcgen05.ld), so every MMA
// accumulates and the block an input row opens needs no separate overwriting MMA: 6 MMAs per row instead of 7 at C = 32, 12 instead of
// 13-14 at C = 64 (one thread gets an N <= 128 MMA through every ~81-94 cycles whatever N is: tools/umma_tput.cu).  Measured -3 % at
// C = 32, -2.4 % at C = 64 (profiles/r02_experiments.md section 8; a second issuing warp in place of a transform warp gained 8 % without
// the GroupNorm prologue and nothing with GroupNorm + SiLU).
#ifndef DDIMB_WIDE_ZERO
#define DDIMB_WIDE_ZERO 1
#endif
constexpr int kWidePos = 130;   // positions per row strip: 128 + 2 halo columns

struct WideParams {
  int32_t B, T, W;
  int32_t RT;             // output rows per work unit
  int32_t units_t;        // ceil(T / RT)
  int32_t num_units;      // B * (W / 128) * units_t
  int32_t prologue;       // 0: x is already activated; 1: GroupNorm affine; 2: GroupNorm affine + SiLU; 3 (C = 32): conv_in -- rows built from src_f32
  const float* src_f32;   // prologue 3: fp32 NCHW network input [B][2][T][W]
  float eps;
  const double* gn_stats; // [B][8][2] (sum, sumsq) of x
  const float* gn_gamma;  // [C]
  const float* gn_beta;   // [C] or null
  // epilogue (same meaning as GemmParams)
  __nv_bfloat16* out;
  const __nv_bfloat16* add;
  const float* bias;
  const float* bias_b;
  double* stats;
  int32_t bias_b_stride;
  int32_t act;
};

template <int C>
struct WideCfg {
  static constexpr int kRowBytes = C * 2;
  static constexpr int kSlotBytes = ((kWidePos * kRowBytes + 1023) / 1024) * 1024;
  // rows in flight (HBM latency x bandwidth needs > 60 KB per SM).  MUST be a multiple of the transform warp count: a transform
  // warp owns rows r = tw (mod kXfWarps), so with kSlots % kXfWarps == 0 it is also the only warp that ever waits on the
  // "raw" barrier of a given slot, and it cannot run a whole phase ahead of it.  With 8 slots and 6 warps a fast warp (affine-only
  // prologue) could reach a slot before the PREVIOUS row of that slot had landed, its parity wait aliased to the phase before, and
  // the CTA deadlocked (seen at T = 4096).
  static constexpr bool kZeroAcc = DDIMB_WIDE_ZERO != 0;
  static constexpr int kIssuers = 1;
  static constexpr int kEpiWarps = 8;                                     // two warps per TMEM lane quarter
  static constexpr int kXfWarps = 6;
  static constexpr int kXfWarp0 = 2 + kEpiWarps;
  static constexpr int kThreads = 32 * (kXfWarp0 + kXfWarps);             // 512: a 17th warp would leave a scheduler's five warps 102 registers each
  static constexpr int kSlots = (C == 32) ? 3 * kXfWarps : kXfWarps;
  static constexpr int kAccBufs = (C == 32) ? 16 : 8;   // output rows in flight in TMEM (the ring wraps inside an input row's three blocks on 2 of kAccBufs rows: those rows cost two MMAs per step)
  static constexpr int kTmemCols = kAccBufs * C;    // 512 columns
  static constexpr int kWBytes = 9 * C * C * 2;
  static constexpr int kSwizzle = (C == 64) ? 2 : 4;
  static constexpr int kSBO = 8 * kRowBytes;
  static constexpr int kXfTabBytes = ((kXfWarps * C * 8 + 1023) / 1024) * 1024;
  static constexpr bool kStage = (C == 32);           // the legacy (skip-add) epilogue's smem-staged stores need 2 KB per warp: C = 32 only
  static constexpr int kStageBytes = 32 * C * 2;      // one epilogue warp's output tile (32 positions x C channels): the TMA store's source
  static constexpr int kStageBufs = (C == 32) ? 2 : 1;  // C = 32 has room for two tiles per warp: a tile is never rewritten while the TMA engine may still read it
  static constexpr int kAuxBytes = 1024 + 4096 + kXfTabBytes + kEpiWarps * kStageBufs * kStageBytes;   // barriers | epilogue bias copies | transform scale/shift copies | store staging
  static constexpr int kSmemBytes = 1024 + kWBytes + kSlots * kSlotBytes + kAuxBytes;
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// unit -> (sample, column block, first output row); units are sample-major so a CTA's samples change rarely
struct WideUnit {
  int b, w0, t0, rows;
  __device__ __forceinline__ void set(const WideParams& p, int u) {
    const int per_b = (p.W >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    w0 = wb << 7;
    rows = min(p.RT, p.T - t0);
  }
};

// Transform warp of conv3x3_wide_kernel: GroupNorm affine (MODE 1) or affine + SiLU (MODE 2) applied in place to the row
// strips this warp owns (every WideCfg<C>::kXfWarps-th).  Lane -> (position p0 + 32/CH * k, 16-byte chunk j): consecutive k are 512 B
// apart for both channel counts.  The loop body is branch-free (mode is a template parameter, only the 2-position tail and the
// two image-edge halo positions are special) so the independent pair chains of a batch interleave: the branchy version spent
// ~1040 warp instructions per row at 6 cycles each, most of them control flow around 16-byte granules.
template <int C, int MODE>
__device__ __forceinline__ void wide_transform_warp(const WideParams& p, uint8_t* a_gen, uint32_t bar_raw, uint32_t bar_ready, float* tab, int tw,
                                                    int lane) {
  using Cfg = WideCfg<C>;
  constexpr int NS = Cfg::kSlots;
  constexpr int CH = C / 8;                           // 16-byte chunks per position
  constexpr int PSTEP = 32 / CH;                      // positions per warp iteration (x row bytes = 512 B)
  constexpr int NFULL = 128 / PSTEP;                  // iterations in which every lane has a position (positions 0..127)
  constexpr int BATCH = (C == 64) ? 4 : 8;
  const int j = lane % CH, p0 = lane / CH;
  // byte offset of this lane's granule in iteration k: p0 row + k * 512, chunk j XOR the row's swizzle phase
  const uint32_t off_e = (uint32_t)(p0 * Cfg::kRowBytes) + (uint32_t)(((C == 64) ? (j ^ (p0 & 7)) : (j ^ ((p0 >> 1) & 3))) << 4);
  const uint32_t off_o = (C == 64) ? (off_e ^ 64u) : off_e;     // C = 64: the phase of rows p0 + 4 (odd k) differs in bit 2 of the chunk index
  const float pre = (MODE == 2) ? 0.5f : 1.0f;                  // silu(a) = u * tanh(u) + u with u = a / 2
  uint64_t sc2[4], sh2[4];
  auto granule = [&](uint4 v, bool zero) -> uint4 {
    uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      uint64_t a2 = ffma2(bf16x2_to_f32x2(w4[e]), sc2[e], sh2[e]);
      if (MODE == 2) a2 = silu2_from_half(a2);
      w4[e] = zero ? 0u : f32x2_to_bf16x2(a2);
    }
    return make_uint4(w4[0], w4[1], w4[2], w4[3]);
  };
  int cur_b = -1;
  uint32_t cnt = 0;
  WideUnit u;
  for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
    u.set(p, unit);
    if (u.b != cur_b) {
      cur_b = u.b;
      __syncwarp();
      for (int c = lane; c < C; c += 32) {
        const int g = c / (C / 8);
        const double n = (double)p.T * p.W * (C / 8);
        const double mean = p.gn_stats[(u.b * 8 + g) * 2] / n;
        double var = p.gn_stats[(u.b * 8 + g) * 2 + 1] / n - mean * mean;
        if (var < 0) var = 0;
        const float rstd = (float)(1.0 / sqrt(var + (double)p.eps));
        const float sc = p.gn_gamma[c] * rstd;
        tab[c] = pre * sc;
        tab[C + c] = pre * ((p.gn_beta ? p.gn_beta[c] : 0.f) - (float)mean * sc);
      }
      __syncwarp();
#pragma unroll
      for (int e = 0; e < 4; ++e) { sc2[e] = pack2(tab[8 * j + 2 * e], tab[8 * j + 2 * e + 1]); sh2[e] = pack2(tab[C + 8 * j + 2 * e], tab[C + 8 * j + 2 * e + 1]); }
    }
    const bool left_edge = (u.w0 == 0), right_edge = (u.w0 + 128 == p.W);
    // first row of this unit owned by this warp, then every WideCfg<C>::kXfWarps-th
    int rr = (int)((uint32_t)(tw + WideCfg<C>::kXfWarps - (int)(cnt % WideCfg<C>::kXfWarps)) % WideCfg<C>::kXfWarps);
    for (; rr < u.rows + 2; rr += WideCfg<C>::kXfWarps) {
      const uint32_t c2 = cnt + (uint32_t)rr;
      const uint32_t slot = c2 % NS, use = c2 / NS;
      const int t = u.t0 - 1 + rr;
      uint8_t* sb = a_gen + slot * Cfg::kSlotBytes;
      mbar_wait(bar_raw + 8 * slot, use & 1, 16);
      if (t >= 0 && t < p.T) {   // a row outside the image is all padding: TMA wrote zeros and affine(0) != 0 must not be applied
        // C = 64 keeps a rolled loop of small batches: there the tensor pipe and the epilogue's MUFU budget bound the kernel and the
        // fully unrolled transform measured 11% SLOWER end to end (it finishes rows in bursts that delay the epilogue warps).
#pragma unroll(C == 64 ? 1 : NFULL / BATCH)
        for (int k0 = 0; k0 < NFULL; k0 += BATCH) {
          uint4 v[BATCH];
#pragma unroll
          for (int i = 0; i < BATCH; ++i) v[i] = *reinterpret_cast<const uint4*>(sb + (((k0 + i) & 1) ? off_o : off_e) + (k0 + i) * 512);
#pragma unroll
          for (int i = 0; i < BATCH; ++i)
            *reinterpret_cast<uint4*>(sb + (((k0 + i) & 1) ? off_o : off_e) + (k0 + i) * 512) =
                granule(v[i], (k0 + i == 0) && left_edge && p0 == 0);
        }
        if (p0 < 2) {   // positions 128, 129 (iteration NFULL; NFULL is even)
          uint4* q = reinterpret_cast<uint4*>(sb + off_e + NFULL * 512);
          *q = granule(*q, right_edge && p0 == 1);
        }
      }
      fence_proxy_async_smem();     // generic-proxy writes -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_ready + 8 * slot);
    }
    cnt += (uint32_t)(u.rows + 2);
  }
}

// conv_in (reference models/diffusion.py:189-198, Conv2d(2 -> 32, k3 p1) of the fp32 NCHW input) as prologue 3 of the C = 32 kernel: the
// "transform" warps BUILD the row strips instead of normalising TMA-delivered ones.  A position's 64-byte row holds
// [x0_hi x1_hi x0_lo x1_lo x0_hi x1_hi 0 0 | 0 ...] (x = hi + lo to 2^-17) against weight rows [w_hi w_hi . w_hi w_hi . w_lo w_lo | 0] per tap
// (ddimb_conv_wimg kind 3), i.e. w.x = w_hi x_hi + w_hi x_lo + w_lo x_hi with fp32 accumulation: fp32-class accuracy from ONE K = 16 step
// per tap; taps and output rows are the wide kernel's row-shifted descriptors and N = 3 C MMAs.  Only the first 16 bytes of a row are ever
// written (the slots are zeroed once when the kernel starts).  (The kernel this replaces built a 128-byte im2col row of all 18 taps per
// position: 9 x the shared-memory writes and conversions, 0.37-0.40 of HBM.)
template <int C>
__device__ __forceinline__ void wide_convert_warp(const WideParams& p, uint8_t* a_gen, uint32_t bar_empty, uint32_t bar_ready, int tw, int lane) {
  using Cfg = WideCfg<C>;
  constexpr int NS = Cfg::kSlots, XF = Cfg::kXfWarps;
  uint32_t cnt = 0;
  WideUnit u;
  for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
    u.set(p, unit);
    int rr = (int)((uint32_t)(tw + XF - (int)(cnt % XF)) % XF);
    for (; rr < u.rows + 2; rr += XF) {
      const uint32_t c2 = cnt + (uint32_t)rr;
      const uint32_t slot = c2 % NS, use = c2 / NS;
      const int t = u.t0 - 1 + rr;
      uint8_t* sb = a_gen + slot * Cfg::kSlotBytes;
      const bool row_in = t >= 0 && t < p.T;
      const float* x0 = p.src_f32 + (((int64_t)u.b * 2) * p.T + (row_in ? t : 0)) * p.W;
      const float* x1 = x0 + (int64_t)p.T * p.W;
      float v0[5], v1[5];
#pragma unroll
      for (int i = 0; i < 5; ++i) {                          // positions lane, lane + 32, ... of the 130 (w = w0 - 1 + position)
        const int pos = lane + 32 * i, w = u.w0 - 1 + pos;
        const bool in = row_in && pos < kWidePos && w >= 0 && w < p.W;
        v0[i] = in ? __ldg(x0 + w) : 0.f;
        v1[i] = in ? __ldg(x1 + w) : 0.f;
      }
      mbar_wait(bar_empty + 8 * slot, (use & 1) ^ 1, 17);    // the MMAs of the row that used this slot have completed
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const int pos = lane + 32 * i;
        if (pos < kWidePos) {
          const __nv_bfloat16 h0 = __float2bfloat16_rn(v0[i]), h1 = __float2bfloat16_rn(v1[i]);
          const __nv_bfloat16 l0 = __float2bfloat16_rn(v0[i] - __bfloat162float(h0)), l1 = __float2bfloat16_rn(v1[i] - __bfloat162float(h1));
          const uint32_t hh = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          const uint32_t ll = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
          *reinterpret_cast<uint4*>(sb + pos * 64 + (((pos >> 1) & 3) << 4)) = make_uint4(hh, ll, hh, 0u);   // logical chunk 0 of a SWIZZLE_64B row
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_ready + 8 * slot);
    }
    cnt += (uint32_t)(u.rows + 2);
  }
}

// Epilogue warp of conv3x3_wide_kernel (no skip operand): TMEM -> registers -> (+bias/temb) -> SiLU -> bf16 -> this warp's
// [32 positions][C] staging tile in the TMA swizzle image (16-byte chunk j of position r at slot j ^ phase(r): the 8 lanes of a
// quarter warp hit 8 different bank groups) -> ONE cp.async.bulk.tensor store per tile.  Everything invariant over the tile loop
// is a template parameter or hoisted; the accumulator buffer is handed back to the MMA warp as soon as it is in registers.
// (ncu r02: the generic epilogue spent ~600 instructions per 32 x 32 tile, most of them address / predicate bookkeeping, and its
// eight warps were the busiest of the CTA.)
template <int C, int ACT, bool STATS>
__device__ __forceinline__ void wide_epilogue_lean(const WideParams& p, const CUtensorMap* tmO, uint32_t tmem_base, uint32_t bar_afull,
                                                   uint32_t bar_aempty, float* wb, uint8_t* stage_gen, int warp, int lane) {
  using Cfg = WideCfg<C>;
  constexpr int NB = Cfg::kAccBufs;
  constexpr uint32_t kGrpMask = WideCfg<C>::kEpiWarps / 4 - 1;
  constexpr int CG = C / 8;
  // C = 32: the warp's 32 positions x 64 B are contiguous in global memory, so they leave as ONE 1-D bulk copy of a linear tile (C = 64:
  // a tensor store of the swizzled tile).  Measured equal to the tensor store at C = 32 (profiles/r02_experiments.md section 1); straight
  // 16-byte stores from registers (no staging tile) measured 3 % (C = 32) and 11 % (C = 64) slower.
  constexpr bool kBulk = (C == 32) && (DDIMB_WIDE_BULKST != 0);

  const int q = warp & 3;
  const uint32_t grp = (uint32_t)(warp - 2) >> 2;
  constexpr int NSB = Cfg::kStageBufs;
  const uint32_t stage0 = smem_u32(stage_gen);
  const uint32_t phase = (C == 64) ? (uint32_t)(lane & 7) : (uint32_t)((lane >> 1) & 3);
  uint32_t nstore = 0;
  const uint64_t s2x = (ACT == 1) ? pack2(0.5f, 0.5f) : pack2(1.0f, 1.0f);
  ConvEpiState<C> st;
  st.reset();
  st.cur_batch = -1;
  uint32_t tile = 0;
  WideUnit u;
  for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
    u.set(p, unit);
    if (u.b != st.cur_batch) {
      if (STATS) st.flush(p.stats, lane);
      st.cur_batch = u.b;
      conv_epilogue_bias<C>(wb, p.bias, p.bias_b, u.b, p.bias_b_stride, ACT, lane);
    }
    const int wq = u.w0 + q * 32;
    for (int ot = (int)((grp - tile) & kGrpMask); ot < u.rows; ot += (int)kGrpMask + 1) {   // this group's output rows only
      const uint32_t buf = (tile + ot) % NB, use = (tile + ot) / NB;
      mbar_wait(bar_afull + 8 * buf, use & 1, 15);
      tc_fence_after();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * C;
      const uint32_t sb = (NSB == 2) ? (nstore & 1u) : 0u;
      const uint32_t stage = stage0 + sb * Cfg::kStageBytes;
      uint8_t* srow = stage_gen + sb * Cfg::kStageBytes + lane * (C * 2);
#pragma unroll
      for (int c0 = 0; c0 < C; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(trow + (uint32_t)c0, v);
        tmem_ld_wait();
        if (c0 + 32 == C) {     // accumulators are in registers: the MMA warp may overwrite the buffer
          if (Cfg::kZeroAcc) {  // ... which it only ever accumulates into: hand it back zeroed
#pragma unroll
            for (int z0 = 0; z0 < C; z0 += 32) tmem_st_zero32(trow + (uint32_t)z0);
            tmem_st_wait();
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
        }
        uint64_t f[16];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(wb + c0 + 2 * j);
          f[j] = ffma2(pack2u(v[2 * j], v[2 * j + 1]), s2x, b4.x);
          f[j + 1] = ffma2(pack2u(v[2 * j + 2], v[2 * j + 3]), s2x, b4.y);
        }
        if (ACT == 1) {
#pragma unroll
          for (int j = 0; j < 16; ++j) f[j] = silu2_from_half(f[j]);
        }
        if (c0 == 0 && nstore >= (uint32_t)NSB) {
          // the staging tile about to be rewritten was the source of the store issued NSB tiles ago: waited for only now, after this
          // tile's tcgen05.ld and arithmetic, so the TMA engine's read overlaps them
          bulk_wait_group_read<NSB - 1>();
          __syncwarp();
        }
        if (kBulk) {
          // linear [32 positions][64 B] tile for a 1-D bulk store: lane writes its chunks in the rotated order (i + phase) & 3 so the 8
          // lanes of a quarter warp still hit 8 different bank groups; the rotation of the four 16-byte register groups is two
          // stages of per-lane selects
          uint4 ch[4];
#pragma unroll
          for (int qq = 0; qq < 4; ++qq)
            ch[qq] = make_uint4(f32x2_to_bf16x2(f[4 * qq]), f32x2_to_bf16x2(f[4 * qq + 1]), f32x2_to_bf16x2(f[4 * qq + 2]), f32x2_to_bf16x2(f[4 * qq + 3]));
          const bool r2 = (phase & 2u) != 0, r1 = (phase & 1u) != 0;
          auto sel = [](bool c, const uint4& a, const uint4& b) { return make_uint4(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z, c ? a.w : b.w); };
          const uint4 a0 = sel(r2, ch[2], ch[0]), a1 = sel(r2, ch[3], ch[1]), a2 = sel(r2, ch[0], ch[2]), a3 = sel(r2, ch[1], ch[3]);
          const uint4 b0 = sel(r1, a1, a0), b1 = sel(r1, a2, a1), b2 = sel(r1, a3, a2), b3 = sel(r1, a0, a3);
          *reinterpret_cast<uint4*>(srow + (((0u + phase) & 3u) << 4)) = b0;
          *reinterpret_cast<uint4*>(srow + (((1u + phase) & 3u) << 4)) = b1;
          *reinterpret_cast<uint4*>(srow + (((2u + phase) & 3u) << 4)) = b2;
          *reinterpret_cast<uint4*>(srow + (((3u + phase) & 3u) << 4)) = b3;
        } else {
#pragma unroll
          for (int qq = 0; qq < 4; ++qq)
            *reinterpret_cast<uint4*>(srow + ((((uint32_t)(c0 >> 3) + qq) ^ phase) << 4)) =
                make_uint4(f32x2_to_bf16x2(f[4 * qq]), f32x2_to_bf16x2(f[4 * qq + 1]), f32x2_to_bf16x2(f[4 * qq + 2]), f32x2_to_bf16x2(f[4 * qq + 3]));
        }
        if (STATS) {
          if (ACT == 3) {
            // training: the PRE-activation is stored; the following GroupNorm normalises silu(stored value)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 t = unpack2(f[j]);
              f[j] = pack2(silu_fast(__bfloat162float(__float2bfloat16_rn(t.x))), silu_fast(__bfloat162float(__float2bfloat16_rn(t.y))));
            }
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int g = (c0 + 2 * j) / CG;   // compile-time after unrolling
            st.s1[g] = fadd2(st.s1[g], f[j]);
            st.s2[g] = ffma2(f[j], f[j], st.s2[g]);
          }
        }
      }
      fence_proxy_async_smem();   // generic-proxy writes of every lane -> visible to the TMA engine
      __syncwarp();
      if (kBulk) bulk_store_s2g_e(p.out + ((((int64_t)u.b * p.T + (u.t0 + ot)) * p.W) + wq) * C, stage, 32 * C * 2);
      else tma_store_4d_e(tmO, stage, 0, wq, u.t0 + ot, u.b);
      bulk_commit_group();
      ++nstore;
    }
    tile += (uint32_t)u.rows;
  }
  if (STATS) st.flush(p.stats, lane);
  bulk_wait_group<0>();
}

// FROM_F32: the conv_in instantiation (prologue 3).  A template parameter, not a run-time mode: as a run-time branch in the issuing
// thread's step loop and the producer it cost every C = 32 call 6 %.
template <int C, bool FROM_F32 = false>
__global__ void __launch_bounds__(WideCfg<C>::kThreads, 1)
conv3x3_wide_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                    const __grid_constant__ WideParams p) {
  using Cfg = WideCfg<C>;
  constexpr int NS = Cfg::kSlots, NB = Cfg::kAccBufs;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t w_base = smem_base;
  const uint32_t a_base = smem_base + Cfg::kWBytes;
  uint8_t* a_gen = smem_gen + Cfg::kWBytes;
  const uint32_t aux = a_base + NS * Cfg::kSlotBytes;
  uint8_t* aux_gen = a_gen + NS * Cfg::kSlotBytes;
  // aux: raw_full[24] @0, ready[24] @192, empty[24] @384, acc_full[16] @576, acc_empty[16] @704, wfull @832, tmem ptr @840
  static_assert(NS <= 24 && NS % WideCfg<C>::kXfWarps == 0, "ring slots: a multiple of the transform warp count");
  const uint32_t bar_raw = aux, bar_ready = aux + 192, bar_empty = aux + 384, bar_afull = aux + 576, bar_aempty = aux + 704, bar_w = aux + 832;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 840);
  float* bias_smem = reinterpret_cast<float*>(aux_gen + 1024);
  float* gn_smem = reinterpret_cast<float*>(aux_gen + 1024 + 4096);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool fused = p.prologue != 0;
  constexpr bool from_f32 = FROM_F32;
  static_assert(!FROM_F32 || C == 32, "conv_in is a C = 32 instantiation");
  if (from_f32) {     // rows are built by the converter warps, 16 bytes per position: everything else in a slot stays zero
    for (int i = threadIdx.x; i < NS * Cfg::kSlotBytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(a_gen)[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_raw + 8 * s, 1);
      mbar_init(bar_ready + 8 * s, 1);    // the transform warp that owns the row
      mbar_init(bar_empty + 8 * s, Cfg::kIssuers);
    }
    for (int b = 0; b < NB; ++b) {
      mbar_init(bar_afull + 8 * b, Cfg::kIssuers);
      mbar_init(bar_aempty + 8 * b, 4);   // one arrive per epilogue warp
    }
    mbar_init(bar_w, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 840, (uint32_t)Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ TMA producer
    if (elect_one()) {
      mbar_arrive_expect_tx(bar_w, (uint32_t)Cfg::kWBytes);
      // per kx one B tile of 3C rows: row block j holds ky = 2 - j, i.e. the output rows (rr-2, rr-1, rr) an input row rr feeds
      for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) tma_load_3d(w_base + (kx * 3 + (2 - ky)) * C * C * 2, &tmB, bar_w, (ky * 3 + kx) * C, 0, 0);
      uint32_t cnt = 0;
      WideUnit u;
      for (int unit = blockIdx.x; unit < p.num_units && !from_f32; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr, ++cnt) {
          const uint32_t slot = cnt % NS, use = cnt / NS;
          mbar_wait(bar_empty + 8 * slot, (use & 1) ^ 1, 11);
          mbar_arrive_expect_tx(bar_raw + 8 * slot, (uint32_t)(kWidePos * Cfg::kRowBytes));
          tma_load_4d(a_base + slot * Cfg::kSlotBytes, &tmA, bar_raw + 8 * slot, 0, u.w0 - 1, u.t0 - 1 + rr, u.b);
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (single thread)
    // Input-row-major issue order: input row rr of a unit feeds output rows rr-2 (ky=2), rr-1 (ky=1) and rr (ky=0), whose
    // accumulators are adjacent TMEM column blocks, so one MMA of N = 3C serves all three and the A tile is read from shared
    // memory once per (kx, k) instead of three times.  Output row rr receives its FIRST contribution here (overwrite).
    // The issuing thread is the critical resource (ncu + tools/umma_tput.cu: the tensor pipe needs 56 cycles per N=96 MMA, a
    // branchy issue path spent ~150): per row the (<= 3) accumulator segments are worked out ONCE, the 3 x C/16 steps are then
    // straight-line code over precomputed descriptor words.
    if (elect_one()) {
      const uint32_t bar_in = fused ? bar_ready : bar_raw;
      mbar_wait(bar_w, 0, 12);
      tc_fence_after();
      constexpr uint32_t kTile16 = (uint32_t)(C * C * 2) >> 4;          // one C x C weight tile, in descriptor units
      const uint64_t desc_hi = make_kmajor_desc(0, Cfg::kSBO, Cfg::kSwizzle);
      const uint32_t w_lo = (w_base & 0x3FFFFu) >> 4;
      uint32_t cnt = 0, otile = 0;
      WideUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr) {
          const uint32_t slot = (cnt + rr) % NS;
          // output blocks j = 0, 1, 2 <-> output rows rr-2, rr-1, rr (ky = 2, 1, 0); block 2 is a first contribution
          const bool v0 = rr >= 2, v1 = rr >= 1 && rr - 1 < u.rows, v2 = rr < u.rows;
          const uint32_t b0 = (otile + rr + NB - 2) % NB, b1 = (b0 + 1) % NB, b2 = (b0 + 2) % NB;
          // The valid blocks fb..lb are a contiguous range; the TMEM ring wraps at most once inside it (where a block's ring
          // index is 0), so the accumulating steps need at most two MMAs: segment A = fb..split-1, segment B = split..lb at column 0.
          const int fb = v0 ? 0 : (v1 ? 1 : 2), lb = v2 ? 2 : (v1 ? 1 : 0);
          int split = lb + 1;
          if (fb < 1 && 1 <= lb && b1 == 0) split = 1;
          else if (fb < 2 && 2 <= lb && b2 == 0) split = 2;
          const bool has_b = split <= lb;
          const uint32_t ta = tmem_base + ((b0 + (uint32_t)fb) % NB) * C;
          const uint32_t wa = (uint32_t)fb * kTile16, wb2 = (uint32_t)split * kTile16;
          const uint32_t ia = make_idesc_bf16(kTileM, (uint32_t)(split - fb) * C), ib = make_idesc_bf16(kTileM, (uint32_t)(lb - split + 1) * C);
          // first step (kx = 0, k = 0): block 2 is overwritten on its own, blocks fb..lbi accumulate
          const int lbi = v2 ? 1 : lb;
          const int spl_i = split < lbi + 1 ? split : lbi + 1;
          const bool has_ai = fb <= lbi, has_bi = split <= lbi;
          const uint32_t iai = make_idesc_bf16(kTileM, (uint32_t)(has_ai ? spl_i - fb : 1) * C);
          const uint32_t ibi = make_idesc_bf16(kTileM, (uint32_t)(has_bi ? lbi - split + 1 : 1) * C);
          mbar_wait(bar_in + 8 * slot, ((cnt + rr) / NS) & 1, 13);
          if (v2) mbar_wait(bar_aempty + 8 * b2, ((((otile + rr) / NB) & 1) ^ 1) ^ (Cfg::kZeroAcc ? 1u : 0u), 14);   // zeroed mode: phase 0 = the initial zeroing
          tc_fence_after();
          const uint32_t a_lo = ((a_base + slot * Cfg::kSlotBytes) & 0x3FFFFu) >> 4;
          if (Cfg::kZeroAcc) {
            // every block fb..lb accumulates in every step
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
              for (int k = 0; k < C / 16; ++k) {
                if (FROM_F32 && k >= 1) continue;            // conv_in: one K = 16 step per tap holds data
                const uint64_t ad = desc_hi | (uint64_t)(a_lo + (uint32_t)(kx * (Cfg::kRowBytes >> 4) + 2 * k));
                const uint32_t wl = w_lo + (uint32_t)(kx * 3) * kTile16 + 2 * k;
                umma_bf16(ta, ad, desc_hi | (uint64_t)(wl + wa), ia, 1u);
                if (has_b) umma_bf16(tmem_base, ad, desc_hi | (uint64_t)(wl + wb2), ib, 1u);
              }
            }
          } else {
            {
              const uint64_t ad = desc_hi | (uint64_t)a_lo;
              if (has_ai) umma_bf16(ta, ad, desc_hi | (uint64_t)(w_lo + wa), iai, 1u);
              if (has_bi) umma_bf16(tmem_base, ad, desc_hi | (uint64_t)(w_lo + wb2), ibi, 1u);
              if (v2) umma_bf16(tmem_base + b2 * C, ad, desc_hi | (uint64_t)(w_lo + 2 * kTile16), make_idesc_bf16(kTileM, (uint32_t)C), 0u);
            }
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
              for (int k = 0; k < C / 16; ++k) {
                if (kx == 0 && k == 0) continue;
                const uint64_t ad = desc_hi | (uint64_t)(a_lo + (uint32_t)(kx * (Cfg::kRowBytes >> 4) + 2 * k));
                const uint32_t wl = w_lo + (uint32_t)(kx * 3) * kTile16 + 2 * k;
                umma_bf16(ta, ad, desc_hi | (uint64_t)(wl + wa), ia, 1u);
                if (has_b) umma_bf16(tmem_base, ad, desc_hi | (uint64_t)(wl + wb2), ib, 1u);
              }
            }
          }
          umma_commit(bar_empty + 8 * slot);                  // this input row is consumed
          if (v0) umma_commit(bar_afull + 8 * b0);            // output row rr-2 is complete
        }
        cnt += u.rows + 2;
        otile += u.rows;
      }
    }
  } else if (warp < 2 + WideCfg<C>::kEpiWarps) {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4; group = output row parity)
    float* wb = bias_smem + (warp - 2) * 128;
    uint8_t* stage = aux_gen + 1024 + 4096 + Cfg::kXfTabBytes + (warp - 2) * Cfg::kStageBufs * Cfg::kStageBytes;
    if (Cfg::kZeroAcc && warp < 6) {
      // every accumulator block starts zeroed; this completes phase 0 of the "empty" barriers the issuers wait on before a block's first use
#pragma unroll
      for (int z0 = 0; z0 < Cfg::kTmemCols; z0 += 32) tmem_st_zero32(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)z0);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0)
        for (int b = 0; b < NB; ++b) mbar_arrive(bar_aempty + 8 * b);
    }
    if (p.add == nullptr) {
      const bool want = p.stats != nullptr;
#define DDIMB_WIDE_EPI_CASE(A)                                                                                        \
  if (p.act == A) {                                                                                                   \
    if (want) wide_epilogue_lean<C, A, true>(p, &tmO, tmem_base, bar_afull, bar_aempty, wb, stage, warp, lane);       \
    else wide_epilogue_lean<C, A, false>(p, &tmO, tmem_base, bar_afull, bar_aempty, wb, stage, warp, lane);           \
  }
      DDIMB_WIDE_EPI_CASE(0) DDIMB_WIDE_EPI_CASE(1) DDIMB_WIDE_EPI_CASE(3)
#undef DDIMB_WIDE_EPI_CASE
    } else {
      // ---- legacy path: skip operand added after the activation (generic tile epilogue, direct / smem-staged stores)
      const int q = warp & 3;
      const uint32_t grp = (uint32_t)(warp - 2) >> 2;
      constexpr uint32_t kGrpMask = WideCfg<C>::kEpiWarps / 4 - 1;
      const int row = q * 32 + lane;
      ConvEpiState<C> st;
      st.reset();
      st.cur_batch = -1;
      uint32_t tile = 0;
      WideUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        if (u.b != st.cur_batch) {
          st.flush(p.stats, lane);
          st.cur_batch = u.b;
          conv_epilogue_bias<C>(wb, p.bias, p.bias_b, u.b, p.bias_b_stride, p.act, lane);
        }
        for (int ot = (int)((grp - tile) & kGrpMask); ot < u.rows; ot += (int)kGrpMask + 1) {   // this group's output rows only
          const uint32_t buf = (tile + ot) % NB, use = (tile + ot) / NB;
          const int64_t addr0 = ((((int64_t)u.b * p.T + (u.t0 + ot)) * p.W) + u.w0 + row) * C;
          mbar_wait(bar_afull + 8 * buf, use & 1, 15);
          tc_fence_after();
          const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * C;
          conv_epilogue_tile<C, C>(trow, wb, p.act, p.add, p.out, addr0, true, p.stats != nullptr, st, Cfg::kStage ? stage : nullptr);
          if (Cfg::kZeroAcc) {
#pragma unroll
            for (int z0 = 0; z0 < C; z0 += 32) tmem_st_zero32(trow + (uint32_t)z0);
            tmem_st_wait();
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
        }
        tile += (uint32_t)u.rows;
      }
      st.flush(p.stats, lane);
    }
  } else if (fused && warp >= Cfg::kXfWarp0) {
    // ================================================================ transform: GroupNorm affine (+SiLU) in place
    float* tab = gn_smem + (warp - Cfg::kXfWarp0) * 2 * C;
    if constexpr (from_f32) wide_convert_warp<C>(p, a_gen, bar_empty, bar_ready, warp - Cfg::kXfWarp0, lane);
    else if (p.prologue == 2) wide_transform_warp<C, 2>(p, a_gen, bar_raw, bar_ready, tab, warp - Cfg::kXfWarp0, lane);
    else wide_transform_warp<C, 1>(p, a_gen, bar_raw, bar_ready, tab, warp - Cfg::kXfWarp0, lane);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)Cfg::kTmemCols);
  }
}

}  // namespace ddimb
