// Weight-gradient GEMM: D[m][n] += sum over rows r of X_tap[r][m] * Y[r][n]  (autograd of nn.Conv2d / nn.ConvTranspose2d /
// nn.Linear on the denoiser path; reference models/diffusion.py:28-40,59-78,105-109,127-128,156 and hf modeling_fnet.py:204-231).
//
// The contraction runs over POSITIONS (rows of the NHWC / row-major tensors), so both operands are channel-contiguous in
// shared memory exactly as TMA delivers them: tcgen05 "MN-major" operands (instruction descriptor bits 15/16), canonical
// layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units -- an atom is [8 positions][64 channels] (SWIZZLE_128B) or
// [8 positions][32 channels] (SWIZZLE_64B), SBO = 8 rows, LBO = distance between channel atoms (tools/mn_probe.cu).
// A job = (tap, M tile of X channels, N tile of Y channels); a CTA owns one job and one split-K slice of the position
// chunks, accumulates in TMEM and adds its fp32 partial into the gradient with red.global.add.
// Warp roles (192 threads): warp0 = TMA producer, warp1 = MMA issuer (+TMEM alloc), warps 2-5 = epilogue.
#pragma once
#include "common.cuh"

namespace ddimb {

constexpr int kTnMaxJobs = 32;
constexpr int kTnThreads = 192;
constexpr int kTnChunk = 128;            // positions per K chunk (8 MMAs of K = 16)
constexpr int kTnXBytes = 128 * 128 * 2;  // X region of a stage: up to 128 channels x 128 positions

struct TnOperand {
  int32_t step[4];   // coordinate advance of dims 1..4 per unit of the driving digit
  int32_t dig[4];    // driving digit of dims 1..4: 0 = w tile, 1 = t tile, 2 = batch, -1 = none
};

struct TnJob {
  int16_t dx[4];     // X coordinate shift of this tap along dims 1..4
  int32_t x_c0;      // first X channel (TMA dim 0) of the M tile
  int32_t y_c0;      // first Y channel of the N tile
  int32_t m_valid;   // rows of D that exist (<= 128)
  int32_t n_valid;   // columns of D that exist (<= 256)
  int64_t out_off;   // element offset of D[0][0]
};

struct TnParams {
  int32_t nchunk[3];   // K-chunk digits: w tiles, t tiles, batch
  int32_t num_chunks;
  int32_t splits;      // split-K slices; grid = njobs * splits
  int32_t njobs;
  int32_t x_at, y_at;  // atom width in channels: 64 (SWIZZLE_128B) or 32 (SWIZZLE_64B)
  int32_t y_bytes;     // Y region of a stage
  int32_t num_stages;
  int32_t tmem_cols;
  TnOperand x, y;
  int64_t stride_m, stride_n;   // element strides of D rows / columns in out
  float* out;
  TnJob jobs[kTnMaxJobs];
};

__device__ __forceinline__ uint64_t make_mn_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t swizzle_code) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)swizzle_code << 61;
  return d;
}

__global__ void __launch_bounds__(kTnThreads, 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmY, const __grid_constant__ TnParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int stage_bytes = kTnXBytes + p.y_bytes;
  const uint32_t aux = smem_base + p.num_stages * stage_bytes;
  const uint32_t bar_full = aux, bar_empty = aux + 64, bar_done = aux + 128;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + p.num_stages * stage_bytes + 160);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int job_i = blockIdx.x % p.njobs, slice = blockIdx.x / p.njobs;
  const TnJob& job = p.jobs[job_i];
  const int m_atoms = (job.m_valid + p.x_at - 1) / p.x_at;
  const int n_atoms = (job.n_valid + p.y_at - 1) / p.y_at;
  const int nmine = slice < p.num_chunks ? (p.num_chunks - slice + p.splits - 1) / p.splits : 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmY);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    mbar_init(bar_done, 1);
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

  if (nmine > 0) {
    if (warp == 0) {
      // ================================================================ TMA producer
      if (elect_one()) {
        const uint32_t x_atom_bytes = (uint32_t)(kTnChunk * p.x_at * 2), y_atom_bytes = (uint32_t)(kTnChunk * p.y_at * 2);
        const uint32_t tx = (uint32_t)m_atoms * x_atom_bytes + (uint32_t)n_atoms * y_atom_bytes;
        int stage = 0;
        uint32_t phase = 0;
        for (int i = 0; i < nmine; ++i) {
          const int chunk = slice + i * p.splits;
          int dgt[3];
          dgt[0] = chunk % p.nchunk[0];
          const int r = chunk / p.nchunk[0];
          dgt[1] = r % p.nchunk[1];
          dgt[2] = r / p.nchunk[1];
          int xc[4], yc[4];
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            xc[d] = (p.x.dig[d] >= 0 ? dgt[p.x.dig[d]] * p.x.step[d] : 0) + job.dx[d];
            yc[d] = (p.y.dig[d] >= 0 ? dgt[p.y.dig[d]] * p.y.step[d] : 0);
          }
          mbar_wait(bar_empty + 8 * stage, phase ^ 1, 21);
          const uint32_t sx = smem_base + stage * stage_bytes, sy = sx + kTnXBytes;
          mbar_arrive_expect_tx(bar_full + 8 * stage, tx);
          for (int a = 0; a < m_atoms; ++a)
            tma_load_5d(sx + a * x_atom_bytes, &tmX, bar_full + 8 * stage, job.x_c0 + a * p.x_at, xc[0], xc[1], xc[2], xc[3]);
          for (int a = 0; a < n_atoms; ++a)
            tma_load_5d(sy + a * y_atom_bytes, &tmY, bar_full + 8 * stage, job.y_c0 + a * p.y_at, yc[0], yc[1], yc[2], yc[3]);
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1) {
      // ================================================================ MMA issuer (one elected thread: common.cuh elect_one)
      if (elect_one()) {
        const uint32_t n_cols = (uint32_t)(n_atoms * p.y_at);
        // kind::f16, D = f32, A = B = bf16, both operands MN-major (bits 15, 16)
        const uint32_t idesc = make_idesc_bf16(128, n_cols) | (1u << 15) | (1u << 16);
        const uint32_t x_row = (uint32_t)p.x_at * 2, y_row = (uint32_t)p.y_at * 2;
        const uint32_t x_code = p.x_at == 64 ? 2u : 4u, y_code = p.y_at == 64 ? 2u : 4u;
        int stage = 0;
        uint32_t phase = 0;
        for (int i = 0; i < nmine; ++i) {
          mbar_wait(bar_full + 8 * stage, phase, 22);
          tc_fence_after();
          const uint32_t sx = smem_base + stage * stage_bytes, sy = sx + kTnXBytes;
#pragma unroll
          for (int k = 0; k < kTnChunk / 16; ++k) {
            const uint64_t ad = make_mn_desc(sx + k * 16 * x_row, kTnChunk * x_row, 8 * x_row, x_code);
            const uint64_t bd = make_mn_desc(sy + k * 16 * y_row, kTnChunk * y_row, 8 * y_row, y_code);
            umma_bf16(tmem_base, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(bar_empty + 8 * stage);
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
        umma_commit(bar_done);
      }
    } else {
      // ================================================================ epilogue: TMEM -> red.global.add.f32
      const int q = warp & 3;
      const int m = q * 32 + lane;
      mbar_wait(bar_done, 0, 23);
      tc_fence_after();
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
      float* op = p.out + job.out_off + (int64_t)m * p.stride_m;
      for (int c0 = 0; c0 < job.n_valid; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(trow + (uint32_t)c0, v);
        tmem_ld_wait();
        if (m < job.m_valid) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (c0 + j < job.n_valid) atomicAdd(op + (int64_t)(c0 + j) * p.stride_n, __uint_as_float(v[j]));
        }
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

}  // namespace ddimb
