// 3x3 convolution at C = 96 (U-Net level 2: W = 64), reference models/diffusion.py:28-40 -- the level where the generic
// implicit-GEMM kernel is L2-bound: it re-reads the activation tile once per tap (9x) and the whole 166 KB weight tensor
// once per 128-position tile (387 KB of L2 traffic per tile against 3 K cycles of tensor work).
//
// Activation-stationary formulation over FLATTENED PADDED positions.  A sample is viewed as T rows of WP = W + 2 positions
// (one zero halo column either side), flat index p = t * WP + (w + 1).  A work unit is 256 consecutive flat positions = two
// 128-row accumulator tiles.  Its input window -- the NR whole padded rows that contain flat positions
// [p0 - WP - 1, p0 + 256 + WP + 1) -- is brought into shared memory ONCE by three TMA boxes (one per 32-channel K chunk;
// out-of-bounds fill supplies every zero of the padding: halo columns, rows -1 and T) and stays there for all nine taps:
// tap (ky, kx) is the same window read through a tcgen05 A descriptor that starts (ky * WP + kx) positions further on (the
// swizzle is a function of the absolute shared-memory address, so a shifted start needs no re-layout).  Weights stream
// through a small ring as pre-swizzled 6 KB tiles (tap, K chunk) by 1-D bulk copies (no per-row TMA cost: ~4 cycles per
// 64-byte box row would exceed the MMA time); every tile feeds both accumulator tiles.  Per 256 positions the kernel moves
// ~89 KB (window) + 166 KB (weights) through L2 instead of 774 KB.  Outputs at halo columns are computed and dropped (2 of 66).
// Warp roles (384 threads): 0 = window producer, 1 = MMA issuer, 2 = weight producer, 3 idle, 4-11 = epilogue (TMEM lane
// quarter = warp % 4; warps 4-7 take accumulator tile 0, warps 8-11 tile 1).
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kWinThreads = 384;
constexpr int kWinC = 96;
constexpr int kWinChunks = 3;                       // K chunks of 32 channels (64-byte rows, SWIZZLE_64B)
constexpr int kWinTile = 128, kWinUnit = 256;       // flat positions per accumulator tile / per work unit
constexpr int kWinBStage = kWinC * 64;              // one (tap, chunk) weight tile: 96 rows x 64 B
constexpr int kWinBStages = 5;
constexpr int kWinAux = 1024 + 4096 + 8 * 2048;     // barriers + tmem pointer | 8 epilogue warps x 128 floats of bias | 8 x 2 KB store staging

struct WinParams {
  int32_t B, T, W, WP;
  int32_t NR;             // padded rows per window
  int32_t plane_bytes;    // NR * WP * 64 rounded up to 1024
  int32_t units_per_b;    // ceil(T * WP / 256)
  int32_t num_units;
  const uint8_t* wswz;    // [9][3][96 x 64 B] pre-swizzled weight tiles
  __nv_bfloat16* out;
  const __nv_bfloat16* add;
  const float* bias;
  const float* bias_b;
  double* stats;
  int32_t bias_b_stride;
  int32_t act;
};

// [Cout][9 * Cin] K-major bf16 -> [tap][chunk][co][64 B] with the SWIZZLE_64B pattern of a 1024-byte aligned tile
// (16-byte chunk index ^= (row >> 1) & 3), i.e. exactly the shared-memory image a TMA load of that tile would produce.
__global__ void __launch_bounds__(256) win_swizzle_weights_kernel(const __nv_bfloat16* __restrict__ w, uint8_t* __restrict__ dst) {
  const int tile = blockIdx.x;                 // tap * 3 + chunk
  const int tap = tile / kWinChunks, chunk = tile % kWinChunks;
  for (int i = threadIdx.x; i < kWinC * 4; i += blockDim.x) {
    const int co = i >> 2, j = i & 3;
    const uint4 v = *reinterpret_cast<const uint4*>(w + (int64_t)co * 9 * kWinC + tap * kWinC + chunk * 32 + j * 8);
    *reinterpret_cast<uint4*>(dst + (int64_t)tile * kWinBStage + co * 64 + ((j ^ ((co >> 1) & 3)) << 4)) = v;
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

__global__ void __launch_bounds__(kWinThreads, 1)
conv3x3_window_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ WinParams p) {
  constexpr int C = kWinC;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t win_bytes = (uint32_t)(kWinChunks * p.plane_bytes);
  const uint32_t a_base = smem_base;                                    // 2 windows x 3 planes
  const uint32_t b_base = a_base + 2 * win_bytes;                       // weight ring
  const uint32_t aux = b_base + kWinBStages * kWinBStage;
  uint8_t* aux_gen = smem_gen + 2 * win_bytes + kWinBStages * kWinBStage;
  // aux: win_full[2] @0, win_empty[2] @16, acc_full[2] @32, acc_empty[2] @48, b_full[] @64, b_empty[] @128, tmem ptr @192
  const uint32_t bar_wfull = aux, bar_wempty = aux + 16, bar_afull = aux + 32, bar_aempty = aux + 48, bar_bfull = aux + 64, bar_bempty = aux + 128;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 192);
  float* bias_smem = reinterpret_cast<float*>(aux_gen + 1024);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_wfull + 8 * i, 1);
      mbar_init(bar_wempty + 8 * i, 1);
      mbar_init(bar_afull + 8 * i, 1);
      mbar_init(bar_aempty + 8 * i, 8);   // one arrive per epilogue warp
    }
    for (int s = 0; s < kWinBStages; ++s) {
      mbar_init(bar_bfull + 8 * s, 1);
      mbar_init(bar_bempty + 8 * s, 1);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 192, 512u);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ window producer: 3 TMA boxes per unit
    if (elect_one()) {
      WinUnit u;
      uint32_t it = 0;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
        u.set(p, unit);
        const uint32_t buf = it & 1;
        mbar_wait(bar_wempty + 8 * buf, ((it >> 1) & 1) ^ 1, 31);
        mbar_arrive_expect_tx(bar_wfull + 8 * buf, (uint32_t)(kWinChunks * p.NR * p.WP * 64));
        for (int ch = 0; ch < kWinChunks; ++ch)
          tma_load_4d(a_base + buf * win_bytes + ch * p.plane_bytes, &tmA, bar_wfull + 8 * buf, ch * 32, -1, u.t_lo, u.b);
      }
    }
  } else if (warp == 2) {
    // ================================================================ weight producer: 27 pre-swizzled tiles per unit
    if (elect_one()) {
      uint32_t cnt = 0;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        for (int tile = 0; tile < 9 * kWinChunks; ++tile, ++cnt) {
          const uint32_t s = cnt % kWinBStages, use = cnt / kWinBStages;
          mbar_wait(bar_bempty + 8 * s, (use & 1) ^ 1, 32);
          mbar_arrive_expect_tx(bar_bfull + 8 * s, (uint32_t)kWinBStage);
          bulk_load_g2s(b_base + s * kWinBStage, p.wswz + (int64_t)tile * kWinBStage, (uint32_t)kWinBStage, bar_bfull + 8 * s);
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread)
    if (elect_one()) {
      const uint64_t desc_hi = make_kmajor_desc(0, 512, 4);              // SWIZZLE_64B, 8-row group pitch 512 B
      const uint32_t idesc = make_idesc_bf16(kWinTile, (uint32_t)C);
      WinUnit u;
      uint32_t it = 0, cnt = 0;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
        u.set(p, unit);
        const uint32_t buf = it & 1, ph = (it >> 1) & 1;
        mbar_wait(bar_aempty + 8 * buf, ph ^ 1, 33);
        mbar_wait(bar_wfull + 8 * buf, ph, 34);
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * (2 * C);
        const uint32_t win = a_base + buf * win_bytes;
        for (int ky = 0; ky < 3; ++ky) {
          for (int kx = 0; kx < 3; ++kx) {
            const uint32_t pos = (uint32_t)(u.off0 + ky * p.WP + kx);
#pragma unroll
            for (int ch = 0; ch < kWinChunks; ++ch, ++cnt) {
              const uint32_t s = cnt % kWinBStages, use = cnt / kWinBStages;
              mbar_wait(bar_bfull + 8 * s, use & 1, 35);
              tc_fence_after();
              const uint32_t a_lo = ((win + ch * p.plane_bytes + pos * 64) & 0x3FFFFu) >> 4;
              const uint32_t b_lo = ((b_base + s * kWinBStage) & 0x3FFFFu) >> 4;
              const uint32_t first = (ky == 0 && kx == 0 && ch == 0) ? 0u : 1u;
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                const uint64_t bd = desc_hi | (uint64_t)(b_lo + 2 * k);
                umma_bf16(acc, desc_hi | (uint64_t)(a_lo + 2 * k), bd, idesc, k ? 1u : first);
                umma_bf16(acc + C, desc_hi | (uint64_t)(a_lo + (kWinTile * 64 >> 4) + 2 * k), bd, idesc, k ? 1u : first);
              }
              umma_commit(bar_bempty + 8 * s);
            }
          }
        }
        umma_commit(bar_wempty + 8 * buf);     // window consumed
        umma_commit(bar_afull + 8 * buf);      // both accumulator tiles complete
      }
    }
  } else if (warp >= 4) {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4, tile = (warp - 4) / 4)
    const int q = warp & 3, g = (warp - 4) >> 2;
    const int m = q * 32 + lane;
    float* wb = bias_smem + (warp - 4) * 128;
    uint8_t* stage = aux_gen + 1024 + 4096 + (warp - 4) * 2048;
    ConvEpiState<C> st;
    st.reset();
    st.cur_batch = -1;
    WinUnit u;
    uint32_t it = 0;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x, ++it) {
      u.set(p, unit);
      if (u.b != st.cur_batch) {
        st.flush(p.stats, lane);
        st.cur_batch = u.b;
        conv_epilogue_bias<C>(wb, p.bias, p.bias_b, u.b, p.bias_b_stride, p.act, lane);
      }
      const uint32_t buf = it & 1, ph = (it >> 1) & 1;
      const int pf = u.p0 + g * kWinTile + m;            // flat padded position of this thread's output
      const int t = pf / p.WP, w = pf - t * p.WP - 1;
      const bool valid = (w >= 0) && (w < p.W) && (t < p.T);
      const int64_t addr0 = ((((int64_t)u.b * p.T + t) * p.W) + w) * C;
      mbar_wait(bar_afull + 8 * buf, ph, 36);
      tc_fence_after();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * (2 * C) + g * C;
      conv_epilogue_tile<C, C>(trow, wb, p.act, p.add, p.out, valid ? addr0 : 0, valid, p.stats != nullptr, st, stage);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
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
