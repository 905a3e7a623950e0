// conv_in on the tensor cores: reference models/diffusion.py:189-198 -- Conv2d(2 -> 32, k3 p1) of the fp32 NCHW network input,
// output bf16 NHWC + GroupNorm partial sums.  K = 18 is no tensor-core shape by itself, and the input is fp32, so four producer
// warps build an im2col row per position in shared memory: 64 bf16 = [x_hi(18) | x_lo(18) | x_hi(18) | 0(10)] (x = hi + lo to
// 2^-17), matched by weight rows [w_hi | w_hi | w_lo | 0]: w.x = w_hi x_hi + w_hi x_lo + w_lo x_hi, fp32-class accuracy from four
// K=16 bf16 MMA steps.  A CTA walks a 128-column strip down the rows; a producer thread owns one column and keeps its
// 2 x 3 x 3 input window in registers, so every input value is loaded once per strip row (6 loads per position).
// The CUDA-core kernel this replaces spent 576 FMAs per position (8.9 ms at the bench shape, 0.17 of HBM).
// Warp roles (512 threads): 0-7 = im2col producers (two sets, alternate rows), 8-15 = epilogue (two groups, alternate rows; TMEM lane
// quarter = warp % 4), the quarter-0 warp of each group issuing the group's MMAs a row ahead.  One warp per role and quarter left both sides latency-bound (4.2 ms vs the figure in DESIGN.md).
#pragma once
#include "conv_wide.cuh"

namespace ddimb {

constexpr int kCiThreads = 512;
constexpr int kCiSlots = 6;                          // im2col rows (128 positions x 128 B) in flight
constexpr int kCiSlotBytes = 128 * 128;
constexpr int kCiNB = 8;                             // output rows in flight in TMEM (32 columns each)
constexpr int kCiWBytes = 32 * 128;                  // [32 co][64 k] bf16, SWIZZLE_128B
constexpr int kCiAux = 1024 + 8192 + 8 * 2048;       // barriers | bias copies | store staging
constexpr int kCiSmem = 1024 + kCiWBytes + kCiSlots * kCiSlotBytes + kCiAux;
constexpr int kCiStripRows = 64;

struct CiParams {
  int32_t B, T, W, RT, units_t, num_units, dbg;
  const float* x;         // fp32 [B][2][T][W]
  const uint8_t* wimg;    // pre-swizzled weight image (kCiWBytes)
  const float* bias;      // [32]
  __nv_bfloat16* out;     // bf16 [B][T][W][32]
  double* stats;
};

// fp32 w[32][2][3][3] -> bf16 [32 rows][64 k]: k < 18 hi(w[k]), 18 <= k < 36 hi(w[k-18]), 36 <= k < 54 lo(w[k-36]), else 0
__global__ void __launch_bounds__(256) ci_weight_image_kernel(const float* __restrict__ w, uint8_t* __restrict__ dst) {
  for (int i = threadIdx.x; i < 32 * 64; i += blockDim.x) {
    const int co = i >> 6, k = i & 63;
    float v = 0.f;
    if (k < 54) {
      const float full = w[co * 18 + (k % 18)];
      const float hi = __bfloat162float(__float2bfloat16_rn(full));
      v = (k < 36) ? hi : full - hi;
    }
    const int j = k >> 3;
    *reinterpret_cast<__nv_bfloat16*>(dst + co * 128 + ((j ^ (co & 7)) << 4) + (k & 7) * 2) = __float2bfloat16_rn(v);
  }
}

struct CiUnit {
  int b, w0, t0, rows;
  __device__ __forceinline__ void set(const CiParams& p, int u) {
    const int per_b = (p.W >> 7) * p.units_t;
    b = u / per_b;
    const int r = u - b * per_b;
    const int wb = r / p.units_t;
    t0 = (r - wb * p.units_t) * p.RT;
    w0 = wb << 7;
    rows = min(p.RT, p.T - t0);
  }
};

__global__ void __launch_bounds__(kCiThreads, 1) conv_in_tc_kernel(const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CiParams p) {
  constexpr int NS = kCiSlots, NB = kCiNB, CO = 32;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t w_base = smem_base;                                   // 4 KB (1024-aligned, so is what follows)
  const uint32_t a_base = smem_base + kCiWBytes;
  uint8_t* a_gen = smem_gen + kCiWBytes;
  const uint32_t aux = a_base + NS * kCiSlotBytes;
  uint8_t* aux_gen = a_gen + NS * kCiSlotBytes;
  // aux: full[8] @0, empty[8] @64, acc_full[8] @128, acc_empty[8] @192, wfull @256, tmem ptr @264
  const uint32_t bar_full = aux, bar_empty = aux + 64, bar_afull = aux + 128, bar_aempty = aux + 192, bar_w = aux + 256;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(aux_gen + 264);
  float* bias_smem = reinterpret_cast<float*>(aux_gen + 1024);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 8 && lane == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 4);      // one arrive per producer warp
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int b = 0; b < NB; ++b) {
      mbar_init(bar_afull + 8 * b, 1);
      mbar_init(bar_aempty + 8 * b, 4);
    }
    mbar_init(bar_w, 1);
    mbar_fence_init();
  }
  if (warp == 8) {
    tmem_alloc(aux + 264, 256u);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp < 8) {
    // ================================================================ im2col producers: thread = column w0 + m; set s takes the rows
    // whose running index has parity s (NS is even, so a set always refills the same slots)
    const int s = warp >> 2, m = threadIdx.x & 127;
    uint32_t base = 0;
    CiUnit u;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      u.set(p, unit);
      const int w = u.w0 + m;
      const float* xb = p.x + (int64_t)u.b * 2 * p.T * p.W;
      const int t_end = min(u.t0 + u.rows, p.T - 1);   // last input row this unit reads
      auto load6 = [&](int tt, float (&d)[2][3]) {
        const bool ok = tt >= 0 && tt <= t_end;
#pragma unroll
        for (int ci = 0; ci < 2; ++ci) {
          const float* xp = xb + ((int64_t)ci * p.T + tt) * p.W;
          d[ci][0] = (ok && w >= 1) ? __ldg(xp + w - 1) : 0.f;
          d[ci][1] = ok ? __ldg(xp + w) : 0.f;
          d[ci][2] = (ok && w + 1 < p.W) ? __ldg(xp + w + 1) : 0.f;
        }
      };
      auto copy6 = [](float (&d)[2][3], const float (&a)[2][3]) {
#pragma unroll
        for (int ci = 0; ci < 2; ++ci)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) d[ci][kx] = a[ci][kx];
      };
      int r = (s - (int)base) & 1;
      float win[3][2][3], pre[2][2][3];   // window rows t-1, t, t+1 as [ky][ci][kx]; pre = the two rows the next iteration adds
      load6(u.t0 + r - 1, win[0]);
      load6(u.t0 + r, win[1]);
      load6(u.t0 + r + 1, win[2]);
      load6(u.t0 + r + 2, pre[0]);
      load6(u.t0 + r + 3, pre[1]);
      for (; r < u.rows; r += 2) {
        float nw[2][2][3];                // issued one iteration (two rows of work) ahead of their use
        load6(u.t0 + r + 4, nw[0]);
        load6(u.t0 + r + 5, nw[1]);
        // 18 values in k order (ci, ky, kx) -> bf16 hi / lo words
        uint32_t hw[9], lw[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) {
          const int k0 = 2 * q, k1 = 2 * q + 1;
          const float a = win[(k0 % 9) / 3][k0 / 9][k0 % 3], b2 = win[(k1 % 9) / 3][k1 / 9][k1 % 3];
          hw[q] = pack_bf16x2(a, b2);
          const float2 h = unpack_bf16x2(hw[q]);
          lw[q] = pack_bf16x2(a - h.x, b2 - h.y);
        }
        const uint32_t cnt = base + (uint32_t)r;
        const uint32_t slot = cnt % NS, use = cnt / NS;
        mbar_wait(bar_empty + 8 * slot, (use & 1) ^ 1, 71);
        uint8_t* dst = a_gen + slot * kCiSlotBytes + m * 128;
        const int sw = m & 7;
        *reinterpret_cast<uint4*>(dst + ((0 ^ sw) << 4)) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
        *reinterpret_cast<uint4*>(dst + ((1 ^ sw) << 4)) = make_uint4(hw[4], hw[5], hw[6], hw[7]);
        *reinterpret_cast<uint4*>(dst + ((2 ^ sw) << 4)) = make_uint4(hw[8], lw[0], lw[1], lw[2]);
        *reinterpret_cast<uint4*>(dst + ((3 ^ sw) << 4)) = make_uint4(lw[3], lw[4], lw[5], lw[6]);
        *reinterpret_cast<uint4*>(dst + ((4 ^ sw) << 4)) = make_uint4(lw[7], lw[8], hw[0], hw[1]);
        *reinterpret_cast<uint4*>(dst + ((5 ^ sw) << 4)) = make_uint4(hw[2], hw[3], hw[4], hw[5]);
        *reinterpret_cast<uint4*>(dst + ((6 ^ sw) << 4)) = make_uint4(hw[6], hw[7], hw[8], 0u);
        *reinterpret_cast<uint4*>(dst + ((7 ^ sw) << 4)) = make_uint4(0u, 0u, 0u, 0u);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_full + 8 * slot);
        copy6(win[0], win[2]);
        copy6(win[1], pre[0]);
        copy6(win[2], pre[1]);
        copy6(pre[0], nw[0]);
        copy6(pre[1], nw[1]);
      }
      base += (uint32_t)u.rows;
    }
  } else {
    // ================================================================ epilogue (TMEM lane quarter = warp % 4; group g takes the rows of
    // parity g, NB is even): + bias, stats, bf16 NHWC.  The quarter-0 warp of each group also issues the group's MMAs (four K=16
    // steps, N = 32), one row ahead of the row it is about to drain -- slot and accumulator depend only on the running row index.
    const int q = warp & 3, g = (warp - 8) >> 2, ew = warp - 8;
    const int m = q * 32 + lane;
    float* wb = bias_smem + ew * 256;
    uint8_t* stage_gen = aux_gen + 1024 + 8192 + ew * 2048;     // [32 positions][64 B] in the SWIZZLE_64B image: the TMA store's source
    const uint32_t stage = smem_u32(stage_gen);
    const uint32_t phase = (uint32_t)((lane >> 1) & 3);
    uint8_t* srow = stage_gen + lane * 64;
    const bool want_stats = p.stats != nullptr;
    bool pending = false;
    uint32_t total = 0;
    CiUnit u;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      u.set(p, unit);
      total += (uint32_t)u.rows;
    }
    const uint64_t desc_hi = make_kmajor_desc(0, 1024, 2);
    const uint32_t w_lo = (w_base & 0x3FFFFu) >> 4;
    const uint32_t idesc = make_idesc_bf16(kTileM, (uint32_t)CO);
    auto issue = [&](uint32_t cnt) {
      if (q == 0 && cnt < total) {
        if (elect_one()) {
          const uint32_t slot = cnt % NS, buf = cnt % NB;
          mbar_wait(bar_full + 8 * slot, (cnt / NS) & 1, 73);
          mbar_wait(bar_aempty + 8 * buf, ((cnt / NB) & 1) ^ 1, 74);
          tc_fence_after();
          const uint32_t a_lo = ((a_base + slot * kCiSlotBytes) & 0x3FFFFu) >> 4;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base + buf * CO, desc_hi | (uint64_t)(a_lo + 2 * k), desc_hi | (uint64_t)(w_lo + 2 * k), idesc, k ? 1u : 0u);
          umma_commit(bar_empty + 8 * slot);
          umma_commit(bar_afull + 8 * buf);
        }
        __syncwarp();
      }
    };
    if (q == 0) {
      if (g == 0 && elect_one()) {
        mbar_arrive_expect_tx(bar_w, (uint32_t)kCiWBytes);
        bulk_load_g2s(w_base, p.wimg, (uint32_t)kCiWBytes, bar_w);
      }
      __syncwarp();
      mbar_wait(bar_w, 0, 72);
      tc_fence_after();
      issue((uint32_t)g);
    }
    ConvEpiState<CO> st;
    st.reset();
    st.cur_batch = -1;
    uint32_t base = 0;
    for (int unit = blockIdx.x; unit < p.num_units; unit += gridDim.x) {
      u.set(p, unit);
      if (u.b != st.cur_batch) {
        st.flush(p.stats, lane);
        st.cur_batch = u.b;
        conv_epilogue_bias<CO>(wb, p.bias, nullptr, u.b, 0, 0, lane);
      }
      for (int r = (g - (int)base) & 1; r < u.rows; r += 2) {
        const uint32_t cnt = base + (uint32_t)r;
        const uint32_t buf = cnt % NB;
        issue(cnt + 2);
        mbar_wait(bar_afull + 8 * buf, (cnt / NB) & 1, 75);
        tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + buf * CO;
        if (pending) {            // the previous tile's store has read the staging tile
          bulk_wait_group_read<0>();
          __syncwarp();
        }
        uint32_t v[32];
        tmem_ld32(trow, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_aempty + 8 * buf);      // accumulators are in registers
        uint64_t f[16];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(wb + 2 * j);
          f[j] = fadd2(pack2u(v[2 * j], v[2 * j + 1]), b4.x);
          f[j + 1] = fadd2(pack2u(v[2 * j + 2], v[2 * j + 3]), b4.y);
        }
#pragma unroll
        for (int qq = 0; qq < 4; ++qq)
          *reinterpret_cast<uint4*>(srow + (((uint32_t)qq ^ phase) << 4)) =
              make_uint4(f32x2_to_bf16x2(f[4 * qq]), f32x2_to_bf16x2(f[4 * qq + 1]), f32x2_to_bf16x2(f[4 * qq + 2]), f32x2_to_bf16x2(f[4 * qq + 3]));
        if (want_stats) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            st.s1[j >> 1] = fadd2(st.s1[j >> 1], f[j]);         // 4 channels per GroupNorm group
            st.s2[j >> 1] = ffma2(f[j], f[j], st.s2[j >> 1]);
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        tma_store_4d_e(&tmO, stage, 0, u.w0 + q * 32, u.t0 + r, u.b);
        bulk_commit_group();
        pending = true;
      }
      base += (uint32_t)u.rows;
    }
    st.flush(p.stats, lane);
    bulk_wait_group<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256u);
  }
}

}  // namespace ddimb
