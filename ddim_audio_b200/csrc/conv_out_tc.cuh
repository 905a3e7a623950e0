// conv_out on the tensor cores: reference models/diffusion.py:284,199-208 -- Conv2d(32 -> 2, k3 p1) of (x + skip), x / skip
// bf16 NHWC, output fp32 NCHW.  conv(x + skip) = conv(x) + conv(skip), so both tensors feed the same accumulators and no sum
// is ever materialised.  Same machinery as csrc/conv_wide.cuh: a CTA walks a 128-column strip down consecutive rows, every
// input row strip (130 positions x 32 channels of x and of skip) is read from HBM once into a ring slot, the taps are
// row-shifted tcgen05 A descriptors, an input row feeds the accumulators of three output rows through ONE MMA of N = 48.
// An accumulator block is 16 columns (the minimum UMMA N): columns 0,1 carry the bf16 high parts of the two output channels'
// weights, columns 2,3 the low parts (w = hi + lo to 2^-17), which the epilogue adds -- fp32-class weights for free, since the
// 14 spare columns exist anyway.  Warp roles (352 threads): 0 = TMA producer, 1 and 10 = MMA issuers, 2-9 = epilogue (two groups, alternate rows).
// The kernel is bound by the rate at which ONE thread can issue small MMAs (~81-94 cycles each whatever N is; 7 per 8 KB input row;
// with the MMAs compiled out it ran at 6.3 TB/s, profiles/r02_experiments.md section 8): two issuing warps take alternate (kx, k) steps.
// Every MMA accumulates (the epilogue hands each accumulator block back zeroed) and each issuer owns its own ring of blocks (TMEM columns
// 0..127 / 128..255), which the epilogue adds: the summation order stays fixed, results are run-to-run identical.
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kCoThreads = 352;                     // TMA producer, MMA issuer, 2 x 4 epilogue warps (alternating output rows), second MMA issuer
constexpr int kCoSlots = 10;                       // input rows in flight (x and skip planes each)
constexpr int kCoPlane = 9216;                     // 130 x 64 B rounded to 1024
constexpr int kCoSlotBytes = 2 * kCoPlane;
constexpr int kCoNB = 8;                           // output rows in flight in TMEM (16 columns each)
constexpr int kCoWBytes = 9 * 16 * 64;             // [kx][ky block][16 rows][64 B], pre-swizzled
constexpr int kCoSmem = 1024 + kCoWBytes + 1024 + kCoSlots * kCoSlotBytes + 1024;

struct CoParams {
  int32_t B, T, W, RT, units_t, num_units, planes;
  const uint8_t* wimg;    // pre-swizzled weight image (kCoWBytes)
  const float* bias;      // [2]
  float* out;             // fp32 [B][2][T][W]
};

// fp32 w[2][32][3][3] -> bf16 tile image: tile (kx, j) holds ky = 2 - j; row r: 0,1 = hi(w[r]), 2,3 = lo(w[r - 2]), 4..15 = 0
__global__ void __launch_bounds__(256) co_weight_image_kernel(const float* __restrict__ w, uint8_t* __restrict__ dst) {
  for (int i = threadIdx.x; i < 9 * 16 * 32; i += blockDim.x) {
    const int ci = i & 31, r = (i >> 5) & 15, tile = i >> 9;
    const int kx = tile / 3, ky = 2 - (tile % 3);
    float v = 0.f;
    if (r < 4) {
      const float full = w[(((r & 1) * 32 + ci) * 3 + ky) * 3 + kx];
      const float hi = __bfloat162float(__float2bfloat16_rn(full));
      v = (r < 2) ? hi : full - hi;
    }
    const int g = ci >> 3;
    *reinterpret_cast<__nv_bfloat16*>(dst + (tile * 16 + r) * 64 + ((g ^ ((r >> 1) & 3)) << 4) + (ci & 7) * 2) = __float2bfloat16_rn(v);
  }
}

struct CoUnit {
  int b, w0, t0, rows;
  __device__ __forceinline__ void set(const CoParams& p, int u) {
    const int per_b = (p.W >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    w0 = wb << 7;
    rows = min(p.RT, p.T - t0);
  }
};

__global__ void __launch_bounds__(kCoThreads, 1)
conv_out_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CoParams p) {
  constexpr int NS = kCoSlots, NB = kCoNB;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t w_base = smem_base;                                   // 9216 B, then 1 KB of padding keeps a_base 1024-aligned
  const uint32_t a_base = smem_base + kCoWBytes + 1024;
  const uint32_t aux = a_base + NS * kCoSlotBytes;
  uint8_t* aux_gen = smem_gen + kCoWBytes + 1024 + NS * kCoSlotBytes;
  // aux: full[16] @0, empty[16] @128, acc_full[8] @256, acc_empty[8] @320, wfull @384, tmem ptr @392
  const uint32_t bar_full = aux, bar_empty = aux + 128, bar_afull = aux + 256, bar_aempty = aux + 320, bar_w = aux + 384;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 392);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmS);
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 2);   // both issuers
    }
    for (int b = 0; b < NB; ++b) {
      mbar_init(bar_afull + 8 * b, 2);
      mbar_init(bar_aempty + 8 * b, 4);
    }
    mbar_init(bar_w, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 392, 256u);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ TMA producer
    if (elect_one()) {
      mbar_arrive_expect_tx(bar_w, (uint32_t)kCoWBytes);
      bulk_load_g2s(w_base, p.wimg, (uint32_t)kCoWBytes, bar_w);
      uint32_t cnt = 0;
      CoUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr, ++cnt) {
          const uint32_t slot = cnt % NS, use = cnt / NS;
          mbar_wait(bar_empty + 8 * slot, (use & 1) ^ 1, 41);
          mbar_arrive_expect_tx(bar_full + 8 * slot, (uint32_t)(p.planes * kWidePos * 64));
          tma_load_4d(a_base + slot * kCoSlotBytes, &tmX, bar_full + 8 * slot, 0, u.w0 - 1, u.t0 - 1 + rr, u.b);
          if (p.planes == 2) tma_load_4d(a_base + slot * kCoSlotBytes + kCoPlane, &tmS, bar_full + 8 * slot, 0, u.w0 - 1, u.t0 - 1 + rr, u.b);
        }
      }
    }
  } else if (warp == 1 || warp == 10) {
    // ================================================================ MMA issuers (one elected thread each; same segment logic as conv_wide.cuh)
    if (elect_one()) {
      const uint32_t me = warp == 1 ? 0u : 1u;
      mbar_wait(bar_w, 0, 42);
      tc_fence_after();
      constexpr uint32_t kTile16 = (16 * 64) >> 4;                       // one 16-row weight block, in descriptor units
      const uint64_t desc_hi = make_kmajor_desc(0, 512, 4);
      const uint32_t w_lo = (w_base & 0x3FFFFu) >> 4;
      uint32_t cnt = 0, otile = 0;
      CoUnit u;
      for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
        u.set(p, unit);
        for (int rr = 0; rr < u.rows + 2; ++rr) {
          const uint32_t slot = (cnt + rr) % NS;
          const bool v0 = rr >= 2, v1 = rr >= 1 && rr - 1 < u.rows, v2 = rr < u.rows;
          const uint32_t b0 = (otile + rr + NB - 2) % NB, b1 = (b0 + 1) % NB, b2 = (b0 + 2) % NB;
          const int fb = v0 ? 0 : (v1 ? 1 : 2), lb = v2 ? 2 : (v1 ? 1 : 0);
          int split = lb + 1;
          if (fb < 1 && 1 <= lb && b1 == 0) split = 1;
          else if (fb < 2 && 2 <= lb && b2 == 0) split = 2;
          const bool has_b = split <= lb;
          const uint32_t tm = tmem_base + me * 128u;                     // this issuer's ring of accumulator blocks
          const uint32_t ta = tm + ((b0 + (uint32_t)fb) % NB) * 16;
          const uint32_t wa = (uint32_t)fb * kTile16, wb2 = (uint32_t)split * kTile16;
          const uint32_t ia = make_idesc_bf16(kTileM, (uint32_t)(split - fb) * 16), ib = make_idesc_bf16(kTileM, (uint32_t)(lb - split + 1) * 16);
          mbar_wait(bar_full + 8 * slot, ((cnt + rr) / NS) & 1, 43);
          if (v2) mbar_wait(bar_aempty + 8 * b2, ((otile + rr) / NB) & 1, 44);   // phase 0 = the initial zeroing
          tc_fence_after();
          const uint32_t a_lo0 = ((a_base + slot * kCoSlotBytes) & 0x3FFFFu) >> 4;
          for (int pl = 0; pl < p.planes; ++pl) {
            const uint32_t a_lo = a_lo0 + (uint32_t)pl * (kCoPlane >> 4);
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                if ((uint32_t)((kx * 2 + k) & 1) != me) continue;
                const uint64_t ad = desc_hi | (uint64_t)(a_lo + (uint32_t)(kx * 4 + 2 * k));
                const uint32_t wl = w_lo + (uint32_t)(kx * 3) * kTile16 + 2 * k;
                umma_bf16(ta, ad, desc_hi | (uint64_t)(wl + wa), ia, 1u);
                if (has_b) umma_bf16(tm, ad, desc_hi | (uint64_t)(wl + wb2), ib, 1u);
              }
            }
          }
          umma_commit(bar_empty + 8 * slot);
          if (v0) umma_commit(bar_afull + 8 * b0);
        }
        cnt += u.rows + 2;
        otile += u.rows;
      }
    }
  } else {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4; the two warp groups take alternate
    // output rows -- one row is a short dependent chain (wait, tcgen05.ld, two stores), so a single group was latency-bound): hi + lo + bias -> fp32 NCHW
    const int q = warp & 3;
    const uint32_t grp = (uint32_t)(warp - 2) >> 2;
    const int m = q * 32 + lane;
    const float bias0 = p.bias[0], bias1 = p.bias[1];
    if (grp == 0) {
      // every accumulator block starts zeroed; the arrivals complete phase 0 of the "empty" barriers
#pragma unroll
      for (int b = 0; b < 2 * NB; ++b) tmem_st_zero16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)b * 16);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0)
        for (int b = 0; b < NB; ++b) mbar_arrive(bar_aempty + 8 * b);
    }
    uint32_t tile0 = 0;
    CoUnit u;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      u.set(p, unit);
      for (int ot = (int)((grp - tile0) & 1u); ot < u.rows; ot += 2) {
        const uint32_t tile = tile0 + (uint32_t)ot;
        const uint32_t buf = tile % NB, use = tile / NB;
        mbar_wait(bar_afull + 8 * buf, use & 1, 45);
        tc_fence_after();
        uint32_t v[16], v1[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 16, v);
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + 128u + buf * 16, v1);
        tmem_ld_wait();
        tmem_st_zero16(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 16);
        tmem_st_zero16(tmem_base + ((uint32_t)(q * 32) << 16) + 128u + buf * 16);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);
        const int64_t o = (((int64_t)u.b * 2) * p.T + (u.t0 + ot)) * p.W + u.w0 + m;
        p.out[o] = (__uint_as_float(v[0]) + __uint_as_float(v1[0])) + (__uint_as_float(v[2]) + __uint_as_float(v1[2])) + bias0;
        p.out[o + (int64_t)p.T * p.W] = (__uint_as_float(v[1]) + __uint_as_float(v1[1])) + (__uint_as_float(v[3]) + __uint_as_float(v1[3])) + bias1;
      }
      tile0 += (uint32_t)u.rows;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256u);
  }
}

}  // namespace ddimb
