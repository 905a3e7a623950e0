// One tcgen05/TMEM/TMA kernel for every dense contraction on the denoiser path
// (reference models/diffusion.py: Conv2d 3x3 :28-40, Conv2d 4x4 s2 :70-78, ConvTranspose2d 4x4 s2 :59-67,
//  nn.Linear :105-109,127-128,156; hf modeling_fnet.py:142-231 Fourier mixing restated as DFT GEMMs).
//
// D[128 rows, n_acc cols] = sum over "taps" of A_tap[128, K_tap] * B_tap[n, K_tap]^T
//   A rows are the 128 positions of a 5-D TMA box (c, d1, d2, d3, d4) of the NHWC/row-major activation; a tap
//   shifts the box (da[]) and picks a channel range -- TMA out-of-bounds zero fill IS the conv zero padding.
//   B rows are output channels of the K-major packed weight (or a per-sample operand for the DFT GEMMs).
// Warp roles (320 threads): warp0 = TMA producer, warp1 = MMA issuer (+TMEM alloc), warps2-9 = epilogue: two warps per TMEM
// lane quarter that split the accumulator columns (TMEM -> registers -> bias/temb/activation/skip-add/GroupNorm partial sums ->
// swizzled shared-memory tile -> TMA store; the skip operand arrives through a TMA load into the same tile).
#pragma once
#include "common.cuh"

namespace ddimb {

constexpr int kMaxTaps = 16;
constexpr int kTileM = 128;
constexpr int kGemmThreads = 320;
constexpr int kGemmEpiWarps = 8;

struct TapDesc {
  int16_t da[4];      // coordinate shift of the A box along dims 1..4
  int32_t a_c0;       // first A channel (dim 0) of this tap
  int32_t b_k0;       // first B k (dim 0) of this tap
  int32_t b_n0;       // B row of segment 0 (before the n-tile offset)
  int16_t nchunk;     // K chunks of CK elements
  int16_t nseg;       // N segments (each n_seg wide: one UMMA + one B box per segment)
  int32_t tmem_col0;  // accumulator column of segment 0
  int32_t img_blk;    // first block of this tap in the pre-swizzled weight image (block = one (chunk, segment) B tile)
  int32_t first_mask; // bit sg: this tap is the first writer of accumulator segment sg's columns in a tile (launch_gemm fills it)
  int32_t k_valid;    // K elements of this tap that exist (0 = all nchunk * CK); the weight image is zero beyond
};

struct GemmParams {
  // ---- tiling
  int32_t box[4];       // A box extents along dims 1..4, product == 128
  int32_t ntile[4];     // tiles along dims 1..4 (dim 4 = batch)
  int32_t num_n_tiles;  // tiles along N (Linear); 1 for convs
  int32_t num_tiles;    // num_n_tiles * prod(ntile)
  int32_t n_seg;        // UMMA N == rows of one B box (multiple of 16, <= 256)
  int32_t n_acc;        // accumulator columns per tile (<= 512)
  int32_t n_tile_stride;  // B row advance per n-tile
  int32_t a_batched;    // A dim-4 coordinate = batch index (else 0)
  int32_t b_batched;    // B dim-2 coordinate = batch index (else 0)
  int32_t b_rows_per_stage;  // max nseg * n_seg
  int32_t num_stages;
  int32_t kps;          // K chunks per pipeline stage (a stage holds up to kps chunks of ONE tap: fewer, larger barrier rounds)
  int32_t tmem_cols;    // power of two >= 32
  int32_t num_acc_bufs; // 1 or 2
  int32_t ntaps;
  TapDesc taps[kMaxTaps];
  // ---- epilogue
  void* out;                // bf16 or fp32
  const __nv_bfloat16* add; // optional bf16 tensor added after the activation, same addressing as out
  const float* bias;        // optional [bias_mod]
  const float* bias_b;      // optional [batch][bias_b_stride] (temb slice)
  double* stats;            // optional [batch][8][2] (sum, sum of squares) of the stored values
  int64_t out_stride[4];    // element strides of dims 1..4 in out/add
  int64_t out_base;         // element offset added to every address
  int64_t out_lo_offset;    // if != 0: also store the bf16 rounding residual at addr + out_lo_offset
  int32_t out_dim[4];       // valid extents of dims 1..4 (rows outside are neither stored nor counted)
  int32_t n_valid;          // valid columns in total (columns >= n_valid are not stored)
  int32_t bias_mod;         // column c uses bias[c % bias_mod]; GroupNorm group = (c % bias_mod) / (bias_mod / 8)
  int32_t bias_b_stride;
  int32_t act;              // 0 none, 1 silu, 2 gelu_new
  int32_t out_f32;          // 1: out is fp32
  // ---- convolution epilogue through TMA (tmO / tmS: 4-D maps {accumulator column, w, t, batch} of out / add)
  int32_t epi_tdim;         // which tile dim (1 or 2) is the row-like one
  // ---- B operand from a pre-swizzled weight image (1-D bulk copies: no per-box-row TMA cost) instead of tmB
  const uint8_t* b_img;     // blocks of n_seg rows x CK k in the shared-memory (swizzled) layout, in consumption order
};

template <int CK>
struct GemmSmem {
  static constexpr int kABytes = kTileM * CK * 2;
  static constexpr int kSwizzle = (CK == 64) ? 2 : 4;  // descriptor code: 2 = 128B, 4 = 64B
  static constexpr int kSBO = 8 * CK * 2;              // 8-row group pitch
  __host__ __device__ static int stage_bytes(int b_rows, int kps) { return kps * (kABytes + b_rows * CK * 2); }
};

constexpr int kGemmAuxBytes = 1024 + kGemmEpiWarps * 1024 + kGemmEpiWarps * 4096;  // barriers, tmem pointer | per-warp bias copies | per-warp [32 x 64 ch] store tiles

// Mixed-radix tile counter (W tiles, T tiles, dim-3 tiles, batch, N tiles) advanced by gridDim.x per step without any
// division: the step is decomposed once, every digit then carries at most 1.
struct TileIter {
  int i[5], n[5], d[5];
  int tile, num_tiles, step;
  __device__ __forceinline__ void init(const GemmParams& p, int first, int stride) {
    n[0] = p.ntile[0]; n[1] = p.ntile[1]; n[2] = p.ntile[2]; n[3] = p.ntile[3]; n[4] = p.num_n_tiles;
    int a = first, b = stride;
#pragma unroll
    for (int k = 0; k < 5; ++k) { i[k] = a % n[k]; a /= n[k]; d[k] = b % n[k]; b /= n[k]; }
    tile = first; num_tiles = p.num_tiles; step = stride;
  }
  __device__ __forceinline__ bool valid() const { return tile < num_tiles; }
  __device__ __forceinline__ void next() {
    tile += step;
    int carry = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      int v = i[k] + d[k] + carry;
      carry = v >= n[k];
      i[k] = carry ? v - n[k] : v;
    }
  }
};

__device__ __forceinline__ void fence_proxy_async_smem_() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 32 lanes x 32 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
        "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

// Epilogue state of one thread (= one accumulator row) of a convolution: GroupNorm partial sums stay in registers across
// tiles (as fp32x2 pairs, so they are fed by FADD2 / FFMA2) and are reduced (warp shuffle, then one double atomicAdd per value
// per warp) only when the sample changes.
template <int CMOD>
struct ConvEpiState {
  uint64_t s1[8], s2[8];   // per group: (sum, sum) and (sum of squares, sum of squares) over even / odd channels
  int cur_batch;
  __device__ __forceinline__ void reset() {
#pragma unroll
    for (int k = 0; k < 8; ++k) { s1[k] = 0ull; s2[k] = 0ull; }
  }
  __device__ __forceinline__ void flush(double* stats, int lane) {
    if (cur_batch < 0 || stats == nullptr) return;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float2 a = unpack2(s1[k]), b = unpack2(s2[k]);
      const float v1 = warp_sum(a.x + a.y), v2 = warp_sum(b.x + b.y);
      if (lane == 0) {
        atomicAdd(stats + (int64_t)cur_batch * 16 + 2 * k, (double)v1);
        atomicAdd(stats + (int64_t)cur_batch * 16 + 2 * k + 1, (double)v2);
      }
    }
    reset();
  }
};

// This warp's shared-memory copy of the per-channel epilogue bias of sample `batch`: (bias + temb slice) * s, where s = 0.5
// when the epilogue applies SiLU (conv_epilogue_tile evaluates it as u*tanh(u)+u on u = (x+b)/2) and 1 otherwise.
template <int CMOD>
__device__ __forceinline__ void conv_epilogue_bias(float* wb, const float* bias, const float* bias_b, int64_t batch, int bias_b_stride, int act, int lane) {
  const float s = (act == 1) ? 0.5f : 1.0f;
  __syncwarp();
  for (int c = lane; c < CMOD; c += 32) wb[c] = s * ((bias ? bias[c] : 0.f) + (bias_b ? bias_b[batch * bias_b_stride + c] : 0.f));
  __syncwarp();
}

// Coalesced request of this warp's `add` (skip) words for one tile: word pre[4 c + i] is granule (lane & 3) of position
// 8 i + lane / 4 in 32-channel chunk c, so four consecutive lanes read the 64 contiguous bytes of one position (every 32-byte
// sector is requested once; a thread reading its own position's 64 B in four 16-byte loads requests each sector twice and
// touches 32 cache lines per instruction).  conv_epilogue_tile redistributes the words through its staging buffer.
template <int NACC>
__device__ __forceinline__ void conv_epilogue_prefetch_add(const __nv_bfloat16* add, int64_t addr0, bool valid, uint4 (&pre)[NACC / 8]) {
  const int lane_ = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int src = 8 * i + (lane_ >> 2);
    const int64_t a = __shfl_sync(0xffffffffu, addr0, src) + (lane_ & 3) * 8;
    const bool ok = __shfl_sync(0xffffffffu, (int)valid, src) != 0;
#pragma unroll
    for (int c = 0; c < NACC / 32; ++c) pre[4 * c + i] = (add != nullptr && ok) ? ldg_nc_v4(add + a + 32 * c) : make_uint4(0u, 0u, 0u, 0u);
  }
}

// One tile of the convolution epilogue: TMEM -> registers -> + (bias + temb) -> activation -> + skip -> bf16 NHWC store,
// GroupNorm partial sums of the stored values.  NACC accumulator columns, column c is channel c % CMOD; wb = this warp's
// shared-memory bias copy (conv_epilogue_bias).  All arithmetic is packed fp32x2: per element 1/2 FFMA2 (bias), 1 MUFU +
// 1/2 FFMA2 (SiLU), 1/2 F2FP (pack), 1/2 FADD2 + 1/2 FFMA2 (statistics).
// `addpre` (optional): the NACC/8 16-byte words of `add` for this thread, loaded by the caller BEFORE it waited for the
// accumulator (conv_epilogue_prefetch_add): the skip tensor does not depend on the MMAs, so its HBM latency is taken off the
// epilogue's critical path (one batch of loads in flight per warp bounded the up-convolutions at ~2 TB/s).
// `stage` (optional): 2 KB of shared memory private to the calling warp.  A thread owns one position, so a direct 16-byte
// store instruction touches 32 different cache lines (one per lane); staged, each 32-channel chunk goes through shared memory
// and is written so that four consecutive lanes cover the 64 contiguous bytes of one position: 8 line segments per
// instruction instead of 32 (the LSU wavefront count, not HBM, bounded the epilogue warps: ncu lg/mio throttle).
template <int NACC, int CMOD>
__device__ __forceinline__ void conv_epilogue_tile(uint32_t trow, const float* wb, int act, const __nv_bfloat16* add, __nv_bfloat16* out,
                                                   int64_t addr0, bool valid, bool want_stats, ConvEpiState<CMOD>& st,
                                                   uint8_t* stage = nullptr, const uint4* addpre = nullptr) {
  constexpr int CG = CMOD / 8;   // channels per GroupNorm group (multiple of 4)
  const float sf = (act == 1) ? 0.5f : 1.0f;
  const uint64_t s2x = pack2(sf, sf);
  const int lane_ = threadIdx.x & 31;
  int64_t saddr[4];
  bool svalid[4];
  if (stage != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {                 // store iteration i: this lane writes granule (lane & 3) of position 8 i + lane / 4
      const int src = 8 * i + (lane_ >> 2);
      saddr[i] = __shfl_sync(0xffffffffu, addr0, src) + (lane_ & 3) * 8;
      svalid[i] = __shfl_sync(0xffffffffu, (int)valid, src) != 0;
    }
  }
#pragma unroll
  for (int c0 = 0; c0 < NACC; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(trow + (uint32_t)c0, v);
    uint4 av[4];
    const bool staged_add = stage != nullptr && (add != nullptr || addpre != nullptr);
    if (staged_add) {
      // coalesced words (prefetched by the caller, or requested here) -> staging -> this thread's own 64 bytes
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (addpre != nullptr) av[i] = addpre[(c0 >> 3) + i];
        else av[i] = svalid[i] ? ldg_nc_v4(add + saddr[i] + c0) : make_uint4(0u, 0u, 0u, 0u);
      }
    } else if (add != nullptr && valid) {
#pragma unroll
      for (int q = 0; q < 4; ++q) av[q] = ldg_nc_v4(add + addr0 + c0 + 8 * q);
    }
    tmem_ld_wait();
    if (staged_add) {
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int pos = 8 * i + (lane_ >> 2);
        *reinterpret_cast<uint4*>(stage + pos * 64 + (((lane_ & 3) ^ ((pos >> 1) & 3)) << 4)) = av[i];
      }
      __syncwarp();
#pragma unroll
      for (int q = 0; q < 4; ++q) av[q] = *reinterpret_cast<const uint4*>(stage + lane_ * 64 + ((q ^ ((lane_ >> 1) & 3)) << 4));
    }
    uint64_t f[16];
#pragma unroll
    for (int j = 0; j < 16; j += 2) {
      const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(wb + ((c0 + 2 * j) % CMOD));
      f[j] = ffma2(pack2u(v[2 * j], v[2 * j + 1]), s2x, b4.x);
      f[j + 1] = ffma2(pack2u(v[2 * j + 2], v[2 * j + 3]), s2x, b4.y);
    }
    if (act == 1) {
#pragma unroll
      for (int j = 0; j < 16; ++j) f[j] = silu2_from_half(f[j]);
    }
    if ((add != nullptr || addpre != nullptr) && valid) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t aw[4] = {av[q].x, av[q].y, av[q].z, av[q].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) f[4 * q + j] = fadd2(f[4 * q + j], bf16x2_to_f32x2(aw[j]));
      }
    }
    if (stage != nullptr) {
      // position-major 64-byte rows, 16-byte granule index XOR (position >> 1) & 3: conflict-free for both access patterns
      __syncwarp();
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<uint4*>(stage + lane_ * 64 + ((q ^ ((lane_ >> 1) & 3)) << 4)) =
            make_uint4(f32x2_to_bf16x2(f[4 * q]), f32x2_to_bf16x2(f[4 * q + 1]), f32x2_to_bf16x2(f[4 * q + 2]), f32x2_to_bf16x2(f[4 * q + 3]));
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int pos = 8 * i + (lane_ >> 2);
        const uint4 v4 = *reinterpret_cast<const uint4*>(stage + pos * 64 + (((lane_ & 3) ^ ((pos >> 1) & 3)) << 4));
        if (svalid[i]) stg_v4(out + saddr[i] + c0, v4);
      }
    }
    if (valid) {
      if (stage == nullptr) {
        __nv_bfloat16* op = out + addr0 + c0;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          stg_v4(op + 8 * q, make_uint4(f32x2_to_bf16x2(f[4 * q]), f32x2_to_bf16x2(f[4 * q + 1]), f32x2_to_bf16x2(f[4 * q + 2]),
                                        f32x2_to_bf16x2(f[4 * q + 3])));
      }
      if (want_stats) {
        if (act == 3) {
          // training: the PRE-activation was stored (its SiLU derivative is needed in backward); the following GroupNorm
          // normalises silu(value), so the statistics are those of silu(value)
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float2 t = unpack2(f[j]);
            f[j] = pack2(silu_fast(__bfloat162float(__float2bfloat16_rn(t.x))), silu_fast(__bfloat162float(__float2bfloat16_rn(t.y))));
          }
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int g = ((c0 + 2 * j) % CMOD) / CG;   // compile-time after unrolling
          st.s1[g] = fadd2(st.s1[g], f[j]);
          st.s2[g] = ffma2(f[j], f[j], st.s2[g]);
        }
      }
    }
  }
}

// GroupNorm partial sums of one 32-channel chunk whose first channel is 32 * G0 (f = 16 packed channel pairs)
template <int CMOD, int G0>
__device__ __forceinline__ void epi_stats_acc(ConvEpiState<CMOD>& st, const uint64_t (&f)[16]) {
  if constexpr (G0 * 32 < CMOD) {
    constexpr int CG = CMOD / 8;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      constexpr int dummy = 0;
      const int g = (G0 * 32 + 2 * j + dummy) / CG;   // compile-time after unrolling
      st.s1[g] = fadd2(st.s1[g], f[j]);
      st.s2[g] = ffma2(f[j], f[j], st.s2[g]);
    }
  }
}

// Convolution epilogue of umma_gemm_kernel for one epilogue warp (TMEM lane quarter q, column half `half`): per tile and per
// CW-channel chunk this warp owns: [TMA load of the skip chunk into the staging tile] -> TMEM -> registers -> (+bias/temb) ->
// activation -> (+skip) -> bf16 -> staging tile in the TMA swizzle image -> ONE cp.async.bulk.tensor store.  Rows outside the
// tensor are clipped by the store (and excluded from the statistics); everything invariant over the loops is hoisted or a template
// parameter.  (ncu r02: the previous per-thread staged-store epilogue cost ~3 K cycles per 32-column chunk -- with four epilogue warps
// every convolution at C >= 96 was epilogue-bound, not tensor-bound.)
template <int NACC, int CMOD, int ACT>
__device__ __forceinline__ void conv_epi_loop_tma(const GemmParams& p, const CUtensorMap* tmO, const CUtensorMap* tmS, uint32_t tmem_base,
                                                  uint32_t bar_tfull, uint32_t bar_tempty, uint32_t bar_skip, float* wb, uint8_t* stage_gen,
                                                  int q, int half, int lane) {
  constexpr int CW = (CMOD % 64 == 0) ? 64 : 32;   // channels per store box: 128-byte (SWIZZLE_128B) or 64-byte (SWIZZLE_64B) rows
  constexpr int NCH = NACC / CW;
  static_assert(NACC % CW == 0 && CW * 2 * 32 <= 4096, "staging tile");
  const bool want_stats = p.stats != nullptr, has_add = p.add != nullptr;
  const uint32_t stage = smem_u32(stage_gen);
  const uint32_t phase = (CW == 64) ? (uint32_t)(lane & 7) : (uint32_t)((lane >> 1) & 3);
  uint8_t* srow = stage_gen + lane * (CW * 2);
  const uint64_t s2x = (ACT == 1) ? pack2(0.5f, 0.5f) : pack2(1.0f, 1.0f);
  // this thread's position inside the tile (validity / statistics) and the warp's sub-box origin (store / skip-load coordinates)
  int r = q * 32 + lane;
  const int r1 = r % p.box[0]; r /= p.box[0];
  const int r2 = r % p.box[1]; r /= p.box[1];
  const int r3 = r % p.box[2]; r /= p.box[2];
  const int r4 = r;
  const int w_off = (q * 32) % p.box[0], t_off = (q * 32) / p.box[0];
  const int tdim = p.epi_tdim;
  ConvEpiState<CMOD> st;
  st.reset();
  st.cur_batch = -1;
  uint32_t skip_ph = 0;
  bool pending = false;
  int local = 0;
  TileIter it;
  for (it.init(p, blockIdx.x, gridDim.x); it.valid(); it.next(), ++local) {
    const int d1 = it.i[0] * p.box[0] + r1, d2 = it.i[1] * p.box[1] + r2, d3 = it.i[2] * p.box[2] + r3, d4 = it.i[3] * p.box[3] + r4;
    const bool valid = d1 < p.out_dim[0] && d2 < p.out_dim[1] && d3 < p.out_dim[2] && d4 < p.out_dim[3];
    const int batch = it.i[3];
    if (batch != st.cur_batch) {
      st.flush(p.stats, lane);
      st.cur_batch = batch;
      conv_epilogue_bias<CMOD>(wb, p.bias, p.bias_b, batch, p.bias_b_stride, ACT, lane);
    }
    const int cw = it.i[0] * p.box[0] + w_off, ct = it.i[tdim] * p.box[tdim] + t_off;
    const int buf = local % p.num_acc_bufs;
    const uint32_t use = (uint32_t)(local / p.num_acc_bufs);
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * NACC);
    bool waited = false;
#pragma unroll 1
    for (int k = half; k < NCH; k += 2) {
      if (pending) {            // the previous store has read the staging tile
        bulk_wait_group_read<0>();
        __syncwarp();
        pending = false;
      }
      if (has_add && elect_one()) {
        mbar_arrive_expect_tx(bar_skip, (uint32_t)(32 * CW * 2));
        asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(stage),
                     "l"(tmS), "r"(bar_skip), "r"(k * CW), "r"(cw), "r"(ct), "r"(batch)
                     : "memory");
      }
      if (!waited) {
        mbar_wait(bar_tfull + 8 * buf, use & 1, 4);
        tc_fence_after();
        waited = true;
      }
#pragma unroll
      for (int c0 = 0; c0 < CW; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(trow + (uint32_t)(k * CW + c0), v);
        tmem_ld_wait();
        if (k + 2 >= NCH && c0 + 32 == CW) {   // this warp's last accumulator columns are in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
        }
        const int col = (k * CW + c0) % CMOD;   // channel of v[0] (bias / GroupNorm group)
        uint64_t f[16];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(wb + col + 2 * j);
          f[j] = ffma2(pack2u(v[2 * j], v[2 * j + 1]), s2x, b4.x);
          f[j + 1] = ffma2(pack2u(v[2 * j + 2], v[2 * j + 3]), s2x, b4.y);
        }
        if (ACT == 1) {
#pragma unroll
          for (int j = 0; j < 16; ++j) f[j] = silu2_from_half(f[j]);
        }
        if (has_add) {
          if (c0 == 0) {
            mbar_wait(bar_skip, skip_ph, 5);
            skip_ph ^= 1u;
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
        if (want_stats && valid) {
          if (ACT == 3) {
            // training: the PRE-activation is stored; the following GroupNorm normalises silu(stored value)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 t = unpack2(f[j]);
              f[j] = pack2(silu_fast(__bfloat162float(__float2bfloat16_rn(t.x))), silu_fast(__bfloat162float(__float2bfloat16_rn(t.y))));
            }
          }
          // st.s1 / st.s2 are register arrays: the group index must be a compile-time constant, so dispatch on the chunk's first
          // channel (a multiple of 32) and let each case unroll with constant indices
          switch (col >> 5) {
            case 0: epi_stats_acc<CMOD, 0>(st, f); break;
            case 1: epi_stats_acc<CMOD, 1>(st, f); break;
            case 2: epi_stats_acc<CMOD, 2>(st, f); break;
            case 3: epi_stats_acc<CMOD, 3>(st, f); break;
            case 4: epi_stats_acc<CMOD, 4>(st, f); break;
            case 5: epi_stats_acc<CMOD, 5>(st, f); break;
            case 6: epi_stats_acc<CMOD, 6>(st, f); break;
            default: epi_stats_acc<CMOD, 7>(st, f); break;
          }
        }
      }
      fence_proxy_async_smem_();   // generic-proxy writes of every lane -> visible to the TMA engine
      __syncwarp();
      tma_store_4d_e(tmO, stage, k * CW, cw, ct, batch);
      bulk_commit_group();
      pending = true;
    }
    if (!waited) {              // fewer chunks than column halves: nothing to do, but the MMA warp counts eight arrivals per buffer
      mbar_wait(bar_tfull + 8 * buf, use & 1, 4);
      if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
    }
  }
  st.flush(p.stats, lane);
  bulk_wait_group<0>();
}

// NACC / CMOD > 0: convolution instantiation (compile-time accumulator width and channel count, lean epilogue with
// GroupNorm partial sums); NACC == 0: generic GEMM epilogue (Linear, Fourier mixing; runtime widths, fp32 or bf16 output).
template <int CK, int NACC, int CMOD>
__global__ void __launch_bounds__(kGemmThreads, 1)
umma_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmO,
                 const __grid_constant__ CUtensorMap tmS, const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // swizzle atoms need 1024-B alignment
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int stage_bytes = GemmSmem<CK>::stage_bytes(p.b_rows_per_stage, p.kps);
  const int b_chunk_bytes = p.b_rows_per_stage * CK * 2;   // B region of one chunk
  const uint32_t aux = smem_base + p.num_stages * stage_bytes;
  // aux layout: full[8] @0, empty[8] @64, tmem_full[2] @128, tmem_empty[2] @144, tmem_ptr @160, skip[8] @192, per-warp bias float[8][256] @1024,
  // per-warp staging tiles [8][4096] @1024 + 8192
  const uint32_t bar_full = aux, bar_empty = aux + 64, bar_tfull = aux + 128, bar_tempty = aux + 144, bar_skip = aux + 192;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + p.num_stages * stage_bytes + 160);
  float* bias_smem = reinterpret_cast<float*>(smem_gen + p.num_stages * stage_bytes + 1024);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_tfull + 8 * b, 1);
      mbar_init(bar_tempty + 8 * b, kGemmEpiWarps);   // one arrive per epilogue warp
    }
    for (int w = 0; w < kGemmEpiWarps; ++w) mbar_init(bar_skip + 8 * w, 1);
    if (NACC > 0) {
      tma_prefetch_desc(&tmO);
      if (p.add != nullptr) tma_prefetch_desc(&tmS);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(aux + 160, (uint32_t)p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ================================================================ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      TileIter it;
      for (it.init(p, blockIdx.x, gridDim.x); it.valid(); it.next()) {
        const int c1 = it.i[0] * p.box[0], c2 = it.i[1] * p.box[1], c3 = it.i[2] * p.box[2];
        const int c4 = p.a_batched ? it.i[3] * p.box[3] : 0;
        const int bb = p.b_batched ? it.i[3] : 0;
        const int nt = it.i[4];
        for (int t = 0; t < p.ntaps; ++t) {
          const TapDesc& tp = p.taps[t];
          const uint32_t tx1 = (uint32_t)(GemmSmem<CK>::kABytes + tp.nseg * p.n_seg * CK * 2);
          for (int ch0 = 0; ch0 < tp.nchunk; ch0 += p.kps) {
            const int nch = min(p.kps, tp.nchunk - ch0);
            mbar_wait(bar_empty + 8 * stage, phase ^ 1, 1);
            const uint32_t sa = smem_base + stage * stage_bytes;
            const uint32_t sb = sa + p.kps * GemmSmem<CK>::kABytes;
            mbar_arrive_expect_tx(bar_full + 8 * stage, tx1 * (uint32_t)nch);
            for (int i = 0; i < nch; ++i) {
              const int ch = ch0 + i;
              tma_load_5d(sa + i * GemmSmem<CK>::kABytes, &tmA, bar_full + 8 * stage, tp.a_c0 + ch * CK, c1 + tp.da[0], c2 + tp.da[1], c3 + tp.da[2],
                          c4 + tp.da[3]);
              if (p.b_img != nullptr) {
                // ncu r02: a TMA box costs ~3.5 cycles per ROW whatever its width; with 18-36 weight boxes of Cout rows per tile the
                // TMA engine, not the tensor pipe, bounded every convolution at C >= 128.  The image is contiguous per block.
                const uint32_t seg_bytes = (uint32_t)(p.n_seg * CK * 2);
                for (int sg = 0; sg < tp.nseg; ++sg)
                  bulk_load_g2s(sb + i * b_chunk_bytes + sg * seg_bytes, p.b_img + (int64_t)(tp.img_blk + ch * tp.nseg + sg) * seg_bytes, seg_bytes,
                                bar_full + 8 * stage);
              } else {
                for (int sg = 0; sg < tp.nseg; ++sg)
                  tma_load_3d(sb + i * b_chunk_bytes + sg * p.n_seg * CK * 2, &tmB, bar_full + 8 * stage, tp.b_k0 + ch * CK,
                              nt * p.n_tile_stride + tp.b_n0 + sg * p.n_seg, bb);
              }
            }
            if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (one elected thread: common.cuh elect_one)
    // ncu r02: this single thread, not the tensor pipe, bounded the convolutions at C >= 128 (~140 instructions per 4-MMA chunk,
    // among them an integer division for the "first write" test).  Which tap writes an accumulator segment first is a static
    // property of the tap program (TapDesc::first_mask); descriptors are built from one constant high word plus 14-bit offsets.
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(kTileM, (uint32_t)p.n_seg);
      const uint64_t desc_hi = make_kmajor_desc(0, GemmSmem<CK>::kSBO, GemmSmem<CK>::kSwizzle);
      const int ntaps = p.ntaps, kps = p.kps, nstages = p.num_stages, nbufs = p.num_acc_bufs, n_acc = p.n_acc;
      const uint32_t seg16 = (uint32_t)(p.n_seg * CK * 2) >> 4, bch16 = (uint32_t)b_chunk_bytes >> 4;
      const uint32_t n_seg = (uint32_t)p.n_seg;
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++local) {
        const int buf = local % nbufs;
        const uint32_t use = (uint32_t)(local / nbufs);
        mbar_wait(bar_tempty + 8 * buf, (use & 1) ^ 1, 2);
        tc_fence_after();
        const uint32_t acc0 = tmem_base + (uint32_t)(buf * n_acc);
        for (int t = 0; t < ntaps; ++t) {
          const int nchunk = p.taps[t].nchunk, nseg = p.taps[t].nseg;
          const uint32_t col0 = acc0 + (uint32_t)p.taps[t].tmem_col0;
          uint32_t first = (uint32_t)p.taps[t].first_mask;     // consumed by the tap's first chunk only
          for (int ch0 = 0; ch0 < nchunk; ch0 += kps) {
            const int nch = min(kps, nchunk - ch0);
            mbar_wait(bar_full + 8 * stage, phase, 3);
            tc_fence_after();
            const uint32_t sa = smem_base + stage * stage_bytes;
            uint32_t a_lo = (sa & 0x3FFFFu) >> 4;
            uint32_t b_lo0 = ((sa + kps * GemmSmem<CK>::kABytes) & 0x3FFFFu) >> 4;
            for (int i = 0; i < nch; ++i, a_lo += (uint32_t)GemmSmem<CK>::kABytes >> 4, b_lo0 += bch16) {
              uint32_t b_lo = b_lo0, col = col0;
              for (int sg = 0; sg < nseg; ++sg, b_lo += seg16, col += n_seg) {
                const uint32_t acc_first = ((first >> sg) & 1u) ^ 1u;
#pragma unroll
                for (int k = 0; k < CK / 16; ++k)   // +32 bytes along K inside the swizzle atom == +2 in the (addr >> 4) start-address field
                  umma_bf16(col, desc_hi | (uint64_t)(a_lo + 2 * k), desc_hi | (uint64_t)(b_lo + 2 * k), idesc, k > 0 ? 1u : acc_first);
              }
              first = 0u;
            }
            umma_commit(bar_empty + 8 * stage);   // smem slot is free once these MMAs have read it
            if (++stage == nstages) { stage = 0; phase ^= 1; }
          }
        }
        umma_commit(bar_tfull + 8 * buf);         // accumulator complete
      }
    }
  } else {
    // ================================================================ epilogue (warps 2..9, TMEM lane quarter = warp % 4, column half = (warp - 2) / 4)
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    int r = row;
    const int r1 = r % p.box[0]; r /= p.box[0];
    const int r2 = r % p.box[1]; r /= p.box[1];
    const int r3 = r % p.box[2]; r /= p.box[2];
    const int r4 = r;
    int local = 0;
    TileIter it;
    if constexpr (NACC > 0) {
      // ---------------- convolution epilogue
      float* wb = bias_smem + (warp - 2) * 256;
      uint8_t* stage = reinterpret_cast<uint8_t*>(bias_smem) + kGemmEpiWarps * 1024 + (warp - 2) * 4096;
      const uint32_t bsk = bar_skip + 8 * (warp - 2);
      if (p.act == 1) conv_epi_loop_tma<NACC, CMOD, 1>(p, &tmO, &tmS, tmem_base, bar_tfull, bar_tempty, bsk, wb, stage, q, half, lane);
      else if (p.act == 3) conv_epi_loop_tma<NACC, CMOD, 3>(p, &tmO, &tmS, tmem_base, bar_tfull, bar_tempty, bsk, wb, stage, q, half, lane);
      else conv_epi_loop_tma<NACC, CMOD, 0>(p, &tmO, &tmS, tmem_base, bar_tfull, bar_tempty, bsk, wb, stage, q, half, lane);
      (void)r1; (void)r2; (void)r3; (void)r4; (void)local;
    } else {
      // ---------------- generic GEMM epilogue (Linear / Fourier mixing)
      for (it.init(p, blockIdx.x, gridDim.x); it.valid(); it.next(), ++local) {
        const int nt = it.i[4];
        const int d1 = it.i[0] * p.box[0] + r1, d2 = it.i[1] * p.box[1] + r2, d3 = it.i[2] * p.box[2] + r3, d4 = it.i[3] * p.box[3] + r4;
        const bool valid = d1 < p.out_dim[0] && d2 < p.out_dim[1] && d3 < p.out_dim[2] && d4 < p.out_dim[3];
        const int64_t addr0 = p.out_base + d1 * p.out_stride[0] + d2 * p.out_stride[1] + d3 * p.out_stride[2] + d4 * p.out_stride[3];
        const int buf = local % p.num_acc_bufs;
        const uint32_t use = (uint32_t)(local / p.num_acc_bufs);
        mbar_wait(bar_tfull + 8 * buf, use & 1, 4);
        tc_fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * p.n_acc);
        for (int c0 = 16 * half; c0 < p.n_acc; c0 += 32) {   // the two warps of a lane quarter alternate 16-column steps
          uint32_t v[16];
          tmem_ld16(trow + (uint32_t)c0, v);
          tmem_ld_wait();
          const int gcol = nt * p.n_acc + c0;   // global column of v[0]
          if (gcol >= p.n_valid) break;         // uniform across the warp
          float f[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
          if (p.bias) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(p.bias + gcol + j);
              f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
            }
          }
          if (p.act == 1) {
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = silu_fast(f[j]);
          } else if (p.act == 2) {
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = gelu_new_fast(f[j]);
          }
          const int nrem = p.n_valid - gcol;    // columns of this chunk that exist (>= 1)
          if (p.out_f32) {
            if (valid) {
              float* op = reinterpret_cast<float*>(p.out) + addr0 + gcol;
              if (nrem >= 16) {
#pragma unroll
                for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(op + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
              } else {
                for (int j = 0; j < nrem; ++j) op[j] = f[j];
              }
            }
          } else {
            uint32_t w[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) w[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
            if (valid) {
              __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.out) + addr0 + gcol;
              if (nrem >= 16) {
                stg_v4(op, make_uint4(w[0], w[1], w[2], w[3]));
                stg_v4(op + 8, make_uint4(w[4], w[5], w[6], w[7]));
              } else {
                for (int j = 0; j < nrem; ++j) op[j] = __float2bfloat16_rn(f[j]);
              }
              if (p.out_lo_offset != 0) {
                // split-bf16: second plane holds bf16(x - bf16(x)) so a later GEMM can use hi*W + lo*W
                uint32_t wl[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float2 h = unpack_bf16x2(w[j]);
                  wl[j] = pack_bf16x2(f[2 * j] - h.x, f[2 * j + 1] - h.y);
                }
                __nv_bfloat16* lp = op + p.out_lo_offset;
                if (nrem >= 16) {
                  stg_v4(lp, make_uint4(wl[0], wl[1], wl[2], wl[3]));
                  stg_v4(lp + 8, make_uint4(wl[4], wl[5], wl[6], wl[7]));
                } else {
                  for (int j = 0; j < nrem; ++j) {
                    const float h = __bfloat162float(__float2bfloat16_rn(f[j]));
                    lp[j] = __float2bfloat16_rn(f[j] - h);
                  }
                }
              }
            }
          }
        }
        // accumulator buffer can be overwritten by the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

// K-major bf16 weight w[row][ldw] -> the blocks of GemmParams' taps in consumption order, each block = n_seg rows x CK k in the
// shared-memory image a TMA load of that box would produce (16-byte chunk j of row r at slot j ^ swizzle phase(r)).
struct WimgParams {
  int32_t ntaps, n_seg, ck, ldw, nblocks;
  TapDesc taps[kMaxTaps];
};
static __global__ void __launch_bounds__(256) wimg_generic_kernel(const __nv_bfloat16* __restrict__ w, uint8_t* __restrict__ img, const __grid_constant__ WimgParams p) {
  const int blk = blockIdx.x;
  int t = 0;
  while (t + 1 < p.ntaps && p.taps[t + 1].img_blk <= blk) ++t;
  const TapDesc& tp = p.taps[t];
  const int rem = blk - tp.img_blk, ch = rem / tp.nseg, sg = rem - ch * tp.nseg;
  const int cpr = p.ck >> 3;                                  // 16-byte chunks per row
  uint8_t* dst = img + (int64_t)blk * p.n_seg * p.ck * 2;
  for (int i = threadIdx.x; i < p.n_seg * cpr; i += blockDim.x) {
    const int row = i / cpr, j = i - row * cpr;
    const int koff = ch * p.ck + j * 8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (tp.k_valid == 0 || koff < tp.k_valid) v = *reinterpret_cast<const uint4*>(w + (int64_t)(tp.b_n0 + sg * p.n_seg + row) * p.ldw + tp.b_k0 + koff);
    const int ph = (p.ck == 64) ? (row & 7) : ((row >> 1) & 3);
    *reinterpret_cast<uint4*>(dst + row * p.ck * 2 + ((j ^ ph) << 4)) = v;
  }
}

}  // namespace ddimb
