// 3x3 convolution at C = 96 (U-Net level 2: W = 64) and C = 128 (level 3: W = 32), reference models/diffusion.py:42-56 --
// [GroupNorm (+SiLU) ->] Conv2d 3x3 -> (+temb | +bias) -> SiLU -- the levels where the generic implicit-GEMM kernel is bound by TMA
// delivery: it re-reads the activation tile once per tap (9x) and the whole weight tensor once per 128-position tile.
//
// Activation-stationary formulation over FLATTENED PADDED positions.  A sample is viewed as T rows of WP = W + 2 positions
// (one zero halo column either side), flat index p = t * WP + (w + 1).  A work unit is 256 consecutive flat positions = two
// 128-row accumulator tiles.  Its input window -- the NR whole padded rows that contain flat positions
// [p0 - WP - 1, p0 + 256 + WP + 1) -- lives in shared memory as C / 32 PLANES of 32 channels (64-byte rows, SWIZZLE_64B; TMA
// out-of-bounds fill supplies every zero of the padding: halo columns, rows -1 and T).  Tap (ky, kx) is the same plane read through
// a tcgen05 A descriptor that starts (ky * WP + kx) positions further on (the swizzle is a function of the absolute shared-memory
// address, so a shifted start needs no re-layout).  Weights stream through a ring as pre-swizzled tiles (channel chunk, tap) of
// C rows x 64 B by 1-D bulk copies; every tile feeds both accumulator tiles.
//   * The window is single-buffered but pipelined PLANE BY PLANE: the MMAs run chunk-major (all nine taps of plane 0, then plane
//     1, ...), so plane c of the next unit is loaded (and normalised) while planes c+1.. of this unit are still being multiplied.
//   * GroupNorm affine (+SiLU) of the block's first half is applied IN PLACE by four transform warps as each plane lands (padding
//     positions stay zero: the reference pads the ACTIVATED tensor), so no activation tensor is materialised between
//     GroupNorm and the convolution (r02: the separate gn_apply passes at these levels were 7 % of the sampling step).
//   * Epilogue: eight lean warps (TMEM lane quarter x accumulator tile), 32-channel chunks -> swizzled staging tile -> ONE TMA tensor
//     store.  The output has no halo columns, so the non-halo positions of a warp's 32 flat positions are CONTIGUOUS in the output
//     viewed as [B][T * W][C]: lanes compact their rows (rank among the non-halo lanes; a warp meets at most one halo pair because
//     WP > 32) and the tile leaves through the tensor map whose box has 32, 31 or 30 rows; the T * W bound clips rows >= T.
//     (TMA stores fault on negative start coordinates, so clipping cannot remove a leading halo position.)
// Per 256 positions the kernel moves the window (~90 KB) + the weights (9 C^2 x 2 B) through L2 instead of 9 A boxes + the weights
// per 128 positions.  Warp roles (512 threads): 0 = window producer, 1 / 3 = MMA issuers of accumulator tile 0 / 1 (a single thread gets one
// N <= 128 MMA through every ~81-94 cycles whatever N is, two threads interleave down to ~60 at N = 96: tools/umma_tput.cu,
// profiles/r02_experiments.md section 8; the tiles share nothing but the operand barriers), 2 = weight producer, 4-7 = epilogue (one
// per TMEM lane quarter, both accumulator tiles), 8-15 = transform (measured: with four transform warps working one 16-byte granule
// at a time the fused kernel was SLOWER than gn_apply + conv; eight warps with four granules in flight each keep the planes ahead
// of the MMA warp).
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kWinThreads = 512;
constexpr int kWinTile = 128, kWinUnit = 256;       // flat positions per accumulator tile / per work unit
constexpr int kWinBStages = 6;
constexpr int kWinEpiWarps = 4, kWinXfWarps = 8;
constexpr int kWinAux = 1024 + kWinEpiWarps * 512 + kWinXfWarps * 1024 + kWinEpiWarps * 2 * 2048;   // barriers | bias copies | scale/shift tables | 2 staging tiles per epilogue warp

template <int C>
struct WinCfg {
  static constexpr int kChunks = C / 32;            // planes of 32 channels
  static constexpr int kBStage = C * 64;            // one (chunk, tap) weight tile: C rows x 64 B
  static constexpr int kWBytes = 9 * kChunks * kBStage;
};

struct WinParams {
  int32_t B, T, W, WP;
  int32_t NR;             // padded rows per window
  int32_t plane_bytes;    // NR * WP * 64 rounded up to 1024
  int32_t units_per_b;    // ceil(T * WP / 256)
  int32_t num_units;
  int32_t prologue;       // 0: x is already activated; 1: GroupNorm affine; 2: GroupNorm affine + SiLU
  float eps;
  const double* gn_stats; // [B][8][2] (sum, sumsq) of x
  const float* gn_gamma;  // [C]
  const float* gn_beta;   // [C] or null
  const uint8_t* wswz;    // [C/32][9][C x 64 B] pre-swizzled weight tiles (chunk-major: the order the MMA warp consumes them)
  const float* bias;
  const float* bias_b;
  double* stats;
  int32_t bias_b_stride;
  int32_t act;
};

// [Cout][9 * Cin] K-major bf16 -> [chunk][tap][co][64 B] with the SWIZZLE_64B pattern of a 1024-byte aligned tile
// (16-byte chunk index ^= (row >> 1) & 3), i.e. exactly the shared-memory image a TMA load of that tile would produce.
template <int C>
__global__ void __launch_bounds__(256) win_swizzle_weights_kernel(const __nv_bfloat16* __restrict__ w, uint8_t* __restrict__ dst) {
  const int tile = blockIdx.x;                 // chunk * 9 + tap
  const int chunk = tile / 9, tap = tile % 9;
  for (int i = threadIdx.x; i < C * 4; i += blockDim.x) {
    const int co = i >> 2, j = i & 3;
    const uint4 v = *reinterpret_cast<const uint4*>(w + (int64_t)co * 9 * C + tap * C + chunk * 32 + j * 8);
    *reinterpret_cast<uint4*>(dst + (int64_t)tile * WinCfg<C>::kBStage + co * 64 + ((j ^ ((co >> 1) & 3)) << 4)) = v;
  }
}

struct WinUnit {
  int b, p0, t_lo, off0;
  __device__ __forceinline__ void set(const WinParams& p, int u) {
    b = u / p.units_per_b;
    p0 = (u - b * p.units_per_b) * kWinUnit;
    // first padded row of the window: floor((p0 - WP - 1) / WP), computed on a non-negative numerator
    t_lo = (p0 + p.WP - 1) / p.WP - 2;
    off0 = p0 - p.WP - 1 - t_lo * p.WP;          // window-relative position of tap (0, 0) of the unit's first output
  }
};

// Epilogue of one accumulator tile for one warp (32 flat padded positions x C channels), see the file comment.
template <int C, int ACT, bool STATS>
__device__ __forceinline__ void win_epilogue_loop(const WinParams& p, const CUtensorMap* tmO32, const CUtensorMap* tmO31, const CUtensorMap* tmO30,
                                                  uint32_t tmem_base, uint32_t bar_afull, uint32_t bar_aempty, float* wb, uint8_t* stage_gen, int q,
                                                  int lane) {
  constexpr int CG = C / 8;
  const uint32_t stage = smem_u32(stage_gen);
  const uint64_t s2x = (ACT == 1) ? pack2(0.5f, 0.5f) : pack2(1.0f, 1.0f);
  ConvEpiState<C> st;
  st.reset();
  st.cur_batch = -1;
  WinUnit u;
  uint32_t it = 0, nstore = 0;
  for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
    u.set(p, unit);
    if (u.b != st.cur_batch) {
      if (STATS) st.flush(p.stats, lane);
      st.cur_batch = u.b;
      conv_epilogue_bias<C>(wb, p.bias, p.bias_b, u.b, p.bias_b_stride, ACT, lane);
    }
    const uint32_t buf = it & 1, ph = (it >> 1) & 1;
    mbar_wait(bar_afull + 8 * buf, ph, 36);
    tc_fence_after();
#pragma unroll 1
    for (int g = 0; g < 2; ++g) {                         // both accumulator tiles of the unit
    const int pf = u.p0 + g * kWinTile + q * 32 + lane;   // this lane's flat padded position
    const int t = pf / p.WP, w = pf - t * p.WP - 1;
    const bool halo = (w < 0) || (w >= p.W);
    const bool valid = !halo && (t < p.T);
    const uint32_t hmask = __ballot_sync(0xffffffffu, halo);
    const int rank = lane - __popc(hmask & ((1u << lane) - 1u));          // row of the compacted staging tile
    const int nrows = 32 - __popc(hmask);                                  // 32, 31 or 30
    const int lin0 = __shfl_sync(0xffffffffu, t * p.W + w, __ffs(~hmask) - 1);   // output position (inside the sample) of the tile's first row
    const CUtensorMap* tmO = nrows == 32 ? tmO32 : (nrows == 31 ? tmO31 : tmO30);
    const uint32_t phase = (uint32_t)((rank >> 1) & 3);
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * (2 * C) + g * C;
#pragma unroll
    for (int c0 = 0; c0 < C; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(trow + (uint32_t)c0, v);
      tmem_ld_wait();
      if (g == 1 && c0 + 32 == C) {                      // accumulators are in registers: the MMA warp may overwrite the buffer
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
      // two staging tiles alternate: the stores issued two chunks ago must have read this one
      const uint32_t sb = nstore & 1u;
      if (nstore >= 2) {
        bulk_wait_group_read<1>();
        __syncwarp();
      }
      uint8_t* srow = stage_gen + sb * 2048 + rank * 64;
      if (!halo) {
#pragma unroll
        for (int qq = 0; qq < 4; ++qq)
          *reinterpret_cast<uint4*>(srow + (((uint32_t)qq ^ phase) << 4)) =
              make_uint4(f32x2_to_bf16x2(f[4 * qq]), f32x2_to_bf16x2(f[4 * qq + 1]), f32x2_to_bf16x2(f[4 * qq + 2]), f32x2_to_bf16x2(f[4 * qq + 3]));
      }
      if (STATS && valid) {
        if (ACT == 3) {
          // training: the PRE-activation is stored; the following GroupNorm normalises silu(stored value)
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float2 tt = unpack2(f[j]);
            f[j] = pack2(silu_fast(__bfloat162float(__float2bfloat16_rn(tt.x))), silu_fast(__bfloat162float(__float2bfloat16_rn(tt.y))));
          }
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int gi = (c0 + 2 * j) / CG;               // compile-time after unrolling
          st.s1[gi] = fadd2(st.s1[gi], f[j]);
          st.s2[gi] = ffma2(f[j], f[j], st.s2[gi]);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      // [C][T * W][B] view of the output: the bound T * W clips the rows of a unit that overhangs the end of the sample
      if (lin0 < p.T * p.W) {
        asm volatile(
            "{\n\t.reg .pred q;\n\t"
            "elect.sync _|q, 0xffffffff;\n\t"
            "@q cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];\n\t}" ::"l"(tmO),
            "r"(stage + sb * 2048), "r"(c0), "r"(lin0), "r"(u.b)
            : "memory");
      }
      bulk_commit_group();
      ++nstore;
    }
    }
  }
  if (STATS) st.flush(p.stats, lane);
  bulk_wait_group<0>();
}

template <int C>
__global__ void __launch_bounds__(kWinThreads, 1)
conv3x3_window_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmO32, const __grid_constant__ CUtensorMap tmO31,
                      const __grid_constant__ CUtensorMap tmO30, const __grid_constant__ WinParams p) {
  using Cfg = WinCfg<C>;
  constexpr int NCH = Cfg::kChunks, NBS = kWinBStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t win_bytes = (uint32_t)(NCH * p.plane_bytes);
  const uint32_t a_base = smem_base;                                    // NCH planes
  const uint32_t b_base = a_base + win_bytes;                           // weight ring
  const uint32_t aux = b_base + NBS * Cfg::kBStage;
  uint8_t* a_gen = smem_gen;
  uint8_t* aux_gen = smem_gen + win_bytes + NBS * Cfg::kBStage;
  // aux: plane_full[4] @0, plane_ready[4] @32, plane_empty[4] @64, acc_full[2] @96, acc_empty[2] @112, b_full[8] @128, b_empty[8] @192, tmem ptr @256
  const uint32_t bar_pfull = aux, bar_pready = aux + 32, bar_pempty = aux + 64, bar_afull = aux + 96, bar_aempty = aux + 112, bar_bfull = aux + 128,
                 bar_bempty = aux + 192;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 256);
  float* bias_smem = reinterpret_cast<float*>(aux_gen + 1024);
  float* gn_smem = reinterpret_cast<float*>(aux_gen + 1024 + kWinEpiWarps * 512);
  uint8_t* stage0 = aux_gen + 1024 + kWinEpiWarps * 512 + kWinXfWarps * 1024;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool fused = p.prologue != 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmO32);
    tma_prefetch_desc(&tmO31);
    tma_prefetch_desc(&tmO30);
    for (int i = 0; i < NCH; ++i) {
      mbar_init(bar_pfull + 8 * i, 1);
      mbar_init(bar_pready + 8 * i, kWinXfWarps);
      mbar_init(bar_pempty + 8 * i, 2);            // both MMA issuers
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_afull + 8 * i, 2);
      mbar_init(bar_aempty + 8 * i, kWinEpiWarps);   // one arrive per epilogue warp
    }
    for (int s = 0; s < NBS; ++s) {
      mbar_init(bar_bfull + 8 * s, 1);
      mbar_init(bar_bempty + 8 * s, 2);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 256, 512u);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ window producer: one TMA box per plane and unit
    if (elect_one()) {
      WinUnit u;
      uint32_t it = 0;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
        u.set(p, unit);
        for (int ch = 0; ch < NCH; ++ch) {
          mbar_wait(bar_pempty + 8 * ch, (it & 1) ^ 1, 31);     // the MMAs of the previous unit have read this plane
          mbar_arrive_expect_tx(bar_pfull + 8 * ch, (uint32_t)(p.NR * p.WP * 64));
          tma_load_4d(a_base + ch * p.plane_bytes, &tmA, bar_pfull + 8 * ch, ch * 32, -1, u.t_lo, u.b);
        }
      }
    }
  } else if (warp == 2) {
    // ================================================================ weight producer: 9 * NCH pre-swizzled tiles per unit, chunk-major
    if (elect_one()) {
      uint32_t cnt = 0;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        for (int tile = 0; tile < 9 * NCH; ++tile, ++cnt) {
          const uint32_t s = cnt % NBS, use = cnt / NBS;
          mbar_wait(bar_bempty + 8 * s, (use & 1) ^ 1, 32);
          mbar_arrive_expect_tx(bar_bfull + 8 * s, (uint32_t)Cfg::kBStage);
          bulk_load_g2s(b_base + s * Cfg::kBStage, p.wswz + (int64_t)tile * Cfg::kBStage, (uint32_t)Cfg::kBStage, bar_bfull + 8 * s);
        }
      }
    }
  } else if (warp == 1 || warp == 3) {
    // ================================================================ MMA issuers (one elected thread each; warp 1 = tile 0, warp 3 = tile 1)
    if (elect_one()) {
      const uint32_t g = warp == 1 ? 0u : 1u;
      const uint64_t desc_hi = make_kmajor_desc(0, 512, 4);              // SWIZZLE_64B, 8-row group pitch 512 B
      const uint32_t idesc = make_idesc_bf16(kWinTile, (uint32_t)C);
      const uint32_t bar_in = fused ? bar_pready : bar_pfull;
      WinUnit u;
      uint32_t it = 0, cnt = 0;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
        u.set(p, unit);
        const uint32_t buf = it & 1, ph = (it >> 1) & 1;
        mbar_wait(bar_aempty + 8 * buf, ph ^ 1, 33);
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * (2 * C) + g * C;
#pragma unroll 1
        for (int ch = 0; ch < NCH; ++ch) {
          mbar_wait(bar_in + 8 * ch, it & 1, 34);
          tc_fence_after();
          const uint32_t plane = a_base + ch * p.plane_bytes;
#pragma unroll 1
          for (int tap = 0; tap < 9; ++tap, ++cnt) {
            const uint32_t pos = (uint32_t)(u.off0 + (tap / 3) * p.WP + (tap % 3));
            const uint32_t s = cnt % NBS, use = cnt / NBS;
            mbar_wait(bar_bfull + 8 * s, use & 1, 35);
            tc_fence_after();
            const uint32_t a_lo = ((plane + (pos + g * kWinTile) * 64) & 0x3FFFFu) >> 4;
            const uint32_t b_lo = ((b_base + s * Cfg::kBStage) & 0x3FFFFu) >> 4;
            const uint32_t first = (ch == 0 && tap == 0) ? 0u : 1u;
#pragma unroll
            for (int k = 0; k < 2; ++k)
              umma_bf16(acc, desc_hi | (uint64_t)(a_lo + 2 * k), desc_hi | (uint64_t)(b_lo + 2 * k), idesc, k ? 1u : first);
            umma_commit(bar_bempty + 8 * s);
          }
          umma_commit(bar_pempty + 8 * ch);      // plane consumed: the next unit's plane may land
        }
        umma_commit(bar_afull + 8 * buf);        // this issuer's accumulator tile is complete
      }
    }
  } else if (warp >= 4 && warp < 4 + kWinEpiWarps) {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4; both tiles of a unit)
    const int q = warp & 3;
    float* wb = bias_smem + (warp - 4) * 128;
    uint8_t* stage = stage0 + (warp - 4) * 4096;
    const bool want = p.stats != nullptr;
#define DDIMB_WIN_EPI_CASE(A)                                                                                      \
  if (p.act == A) {                                                                                                \
    if (want) win_epilogue_loop<C, A, true>(p, &tmO32, &tmO31, &tmO30, tmem_base, bar_afull, bar_aempty, wb, stage, q, lane);     \
    else win_epilogue_loop<C, A, false>(p, &tmO32, &tmO31, &tmO30, tmem_base, bar_afull, bar_aempty, wb, stage, q, lane);         \
  }
    DDIMB_WIN_EPI_CASE(0) DDIMB_WIN_EPI_CASE(1) DDIMB_WIN_EPI_CASE(3)
#undef DDIMB_WIN_EPI_CASE
  } else if (warp >= 4 + kWinEpiWarps && fused) {
    // ================================================================ transform: GroupNorm affine (+SiLU) in place, plane by plane
    const int tw = warp - (4 + kWinEpiWarps);
    float* tab = gn_smem + tw * 256;                     // [C] scale | [C] shift of the current sample
    const int j = lane & 3, pl = lane >> 2;              // 16-byte chunk of the 64-byte row, position inside the 8-position group
    const bool silu = p.prologue == 2;
    const float pre = silu ? 0.5f : 1.0f;                // silu(a) = u * tanh(u) + u with u = a / 2
    const int npos = p.NR * p.WP;
    int cur_b = -1;
    WinUnit u;
    uint32_t it = 0;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
      u.set(p, unit);
      if (u.b != cur_b) {
        cur_b = u.b;
        __syncwarp();
        for (int c = lane; c < C; c += 32) {
          const int gi = c / (C / 8);
          const double n = (double)p.T * p.W * (C / 8);
          const double mean = p.gn_stats[(u.b * 8 + gi) * 2] / n;
          double var = p.gn_stats[(u.b * 8 + gi) * 2 + 1] / n - mean * mean;
          if (var < 0) var = 0;
          const float rstd = (float)(1.0 / sqrt(var + (double)p.eps));
          const float sc = p.gn_gamma[c] * rstd;
          tab[c] = pre * sc;
          tab[C + c] = pre * ((p.gn_beta ? p.gn_beta[c] : 0.f) - (float)mean * sc);
        }
        __syncwarp();
      }
      for (int ch = 0; ch < NCH; ++ch) {
        uint64_t sc2[4], sh2[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int c = ch * 32 + 8 * j + 2 * e;
          sc2[e] = pack2(tab[c], tab[c + 1]);
          sh2[e] = pack2(tab[C + c], tab[C + c + 1]);
        }
        uint8_t* plane = a_gen + ch * p.plane_bytes;
        mbar_wait(bar_pfull + 8 * ch, it & 1, 37);
        // this warp's blocks of 32 positions (four 8-position groups = four independent 16-byte granules per lane in flight)
        for (int g0 = tw * 32; g0 < npos; g0 += kWinXfWarps * 32) {
          uint4 v[4];
          bool ok[4];
          int pos = g0 + pl;
          int r = pos / p.WP, c = pos - r * p.WP;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int t = u.t_lo + r;
            // padding (halo columns, rows outside the image) was zero-filled by TMA and must stay zero
            ok[k] = pos + 8 * k < npos && c >= 1 && c <= p.W && t >= 0 && t < p.T;
            if (ok[k]) v[k] = *reinterpret_cast<const uint4*>(plane + (pos + 8 * k) * 64 + ((j ^ (((pos + 8 * k) >> 1) & 3)) << 4));
            c += 8;
            if (c >= p.WP) { c -= p.WP; ++r; }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (ok[k]) {
              uint32_t w4[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                uint64_t a2 = ffma2(bf16x2_to_f32x2(w4[e]), sc2[e], sh2[e]);
                if (silu) a2 = silu2_from_half(a2);
                w4[e] = f32x2_to_bf16x2(a2);
              }
              *reinterpret_cast<uint4*>(plane + (pos + 8 * k) * 64 + ((j ^ (((pos + 8 * k) >> 1) & 3)) << 4)) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
            }
          }
        }
        fence_proxy_async_smem();     // generic-proxy writes -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_pready + 8 * ch);
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
