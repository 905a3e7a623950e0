// Host side of the weight-gradient GEMM (csrc/gemm_tn.cuh): tensor maps, job tables, C-ABI entry points.
// Every entry point ACCUMULATES into a caller-zeroed fp32 gradient laid out exactly like the torch parameter.
#include <string.h>

#include "../../include/ddim_b200.h"
#include "gemm_tn.cuh"
#include "wgrad_wide.cuh"
#include "wgrad_updown_wide.cuh"
#include "tmap.cuh"

using namespace ddimb;

// ctas_per_sm: split-K CTAs per SM over all jobs.  Every split adds its whole [M x N] tile to the gradient with scattered atomics (tap stride
// 9 floats), so more splits trade shorter K loops for more atomic traffic: measured on the 3x3 gradients at C = 96 .. 256, one CTA per SM
// beats two by 5-37 % and half a CTA per SM is slower again (profiles/r02_experiments.md section 10).
static int launch_tn(TnParams& p, const CUtensorMap& tmX, const CUtensorMap& tmY, cudaStream_t stream, int ctas_per_sm = 2) {
  DDIMB_CHECK_ARG(p.njobs >= 1 && p.njobs <= kTnMaxJobs);
  DDIMB_CHECK_ARG((p.x_at == 32 || p.x_at == 64) && (p.y_at == 32 || p.y_at == 64));
  int max_n = 0;
  for (int j = 0; j < p.njobs; ++j) {
    const TnJob& jb = p.jobs[j];
    DDIMB_CHECK_ARG(jb.m_valid >= 1 && jb.m_valid <= 128 && jb.n_valid >= 1 && jb.n_valid <= 256);
    DDIMB_CHECK_ARG(jb.n_valid % 16 == 0);
    const int n_cols = ceil_div(jb.n_valid, p.y_at) * p.y_at;
    if (n_cols > max_n) max_n = n_cols;
  }
  p.y_bytes = max_n * kTnChunk * 2;
  p.num_chunks = p.nchunk[0] * p.nchunk[1] * p.nchunk[2];
  DDIMB_CHECK_ARG(p.num_chunks >= 1);
  int splits = (ctas_per_sm * ddimb_num_sms()) / p.njobs;
  if (splits < 1) splits = 1;
  if (splits > p.num_chunks) splits = p.num_chunks;
  p.splits = splits;
  p.tmem_cols = next_pow2_cols(max_n);
  const int stage_bytes = kTnXBytes + p.y_bytes;
  const int max_smem = 232448;
  int stages = (max_smem - 1024 - 1024) / stage_bytes;
  if (stages > 6) stages = 6;
  DDIMB_CHECK_ARG(stages >= 2);
  p.num_stages = stages;
  int smem = 1024 + stages * stage_bytes + 1024;
  if (smem < 120 * 1024) smem = 120 * 1024;   // one CTA per SM: TMEM allocation never contends
  DDIMB_SMEM_ATTR(gemm_tn_kernel, max_smem);
  gemm_tn_kernel<<<p.njobs * p.splits, kTnThreads, smem, stream>>>(tmX, tmY, p);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

static void tn_init(TnParams& p) {
  memset(&p, 0, sizeof(p));
  for (int d = 0; d < 4; ++d) { p.x.dig[d] = -1; p.y.dig[d] = -1; }
  p.nchunk[0] = p.nchunk[1] = p.nchunk[2] = 1;
}

static inline int atom_of(int c) { return (c % 64 == 0) ? 64 : 32; }

static void tile_128(int W, int* bw, int* bt) {
  *bw = W >= kTnChunk ? kTnChunk : W;
  *bt = kTnChunk / *bw;
}

// 5-D map over an NHWC tensor [B][T][W][C] as (C, W, T, 1, B) with box (at, bw, bt, 1, 1)
static int nhwc_map(CUtensorMap* tm, const void* ptr, int B, int T, int W, int C, int at, int bw, int bt) {
  uint64_t dims[5] = {(uint64_t)C, (uint64_t)W, (uint64_t)T, 1, (uint64_t)B};
  uint64_t str[4] = {(uint64_t)C, (uint64_t)W * C, (uint64_t)T * W * C, (uint64_t)T * W * C};
  uint32_t box[5] = {(uint32_t)at, (uint32_t)bw, (uint32_t)bt, 1, 1};
  return make_tmap_bf16(tm, ptr, 5, dims, str, box, at);
}
static void nhwc_digits(TnOperand& o, int bw, int bt) {
  o.dig[0] = 0; o.step[0] = bw;
  o.dig[1] = 1; o.step[1] = bt;
  o.dig[3] = 2; o.step[3] = 1;
}

// 5-D space-to-depth map over a high-resolution NHWC tensor [B][T][W][C] as (2C, W/2, 2, T/2, B), box (at, bw, 1, bt, 1):
// the operand of a stride-2 4x4 convolution (csrc/gemm_host.cu ddimb_conv_down4x4)
static int s2d_map(CUtensorMap* tm, const void* ptr, int B, int T, int W, int C, int at, int bw, int bt) {
  uint64_t dims[5] = {(uint64_t)2 * C, (uint64_t)W / 2, 2, (uint64_t)T / 2, (uint64_t)B};
  uint64_t str[4] = {(uint64_t)2 * C, (uint64_t)W * C, (uint64_t)2 * W * C, (uint64_t)T * W * C};
  uint32_t box[5] = {(uint32_t)at, (uint32_t)bw, 1, (uint32_t)bt, 1};
  return make_tmap_bf16(tm, ptr, 5, dims, str, box, at);
}
static void s2d_digits(TnOperand& o, int bw, int bt) {
  o.dig[0] = 0; o.step[0] = bw;
  o.dig[2] = 1; o.step[2] = bt;
  o.dig[3] = 2; o.step[3] = 1;
}

// ------------------------------------------------------------------------------------------------ Conv2d 3x3 weight gradient
// dw[co][ci][ky][kx] += sum_{b,t,w} dy[b,t,w,co] * x[b,t+ky-1,w+kx-1,ci]   (autograd of models/diffusion.py:28-40)
extern "C" int ddimb_conv3x3_wgrad(const void* x, const void* dy, float* dw, int B, int T, int W, int Cin, int Cout, cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && dy && dw && B >= 1 && T >= 1);
  const double npos = (double)B * T * W;
  DDIMB_PROF(stream, "wgrad3x3", Cin, W, npos * (Cin + Cout) * 2 + 36.0 * Cin * Cout, 18.0 * Cin * Cout * npos);
  DDIMB_CHECK_ARG(W >= 8 && W <= 256 && (W & (W - 1)) == 0 && Cin % 32 == 0 && Cout % 32 == 0 && Cin <= 256 && Cout <= 256);
  if (Cin == Cout && (Cin == 32 || Cin == 64) && W % 128 == 0) {
    // two widest levels: x and dy are read once, the accumulators of all nine taps stay in tensor memory (csrc/wgrad_wide.cuh)
    WgParams q;
    memset(&q, 0, sizeof(q));
    q.B = B; q.T = T; q.W = W; q.dw = dw;
    int rt = 64;
    while (rt > 8 && (long long)B * (W / 128) * ceil_div(T, rt) < 2LL * ddimb_num_sms()) rt /= 2;
    q.RT = rt; q.units_t = ceil_div(T, rt); q.num_units = B * (W / 128) * q.units_t;
    CUtensorMap tmX, tmY;
    uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)T, (uint64_t)B};
    uint64_t str[3] = {(uint64_t)Cin, (uint64_t)W * Cin, (uint64_t)T * W * Cin};
    uint32_t bx[4] = {(uint32_t)Cin, (uint32_t)kWgXPos, 1, 1}, by[4] = {(uint32_t)Cin, 128, 1, 1};
    int r = make_tmap_bf16(&tmX, x, 4, dims, str, bx, Cin);
    if (r) return r;
    r = make_tmap_bf16(&tmY, dy, 4, dims, str, by, Cin);
    if (r) return r;
    const int grid = q.num_units < ddimb_num_sms() ? q.num_units : ddimb_num_sms();
    if (Cin == 32) {
      DDIMB_SMEM_ATTR(wgrad_wide_kernel<32>, WgCfg<32>::kSmem);
      wgrad_wide_kernel<32><<<grid, kWgThreads, WgCfg<32>::kSmem, stream>>>(tmX, tmY, q);
    } else {
      DDIMB_SMEM_ATTR(wgrad_wide_kernel<64>, WgCfg<64>::kSmem);
      wgrad_wide_kernel<64><<<grid, kWgThreads, WgCfg<64>::kSmem, stream>>>(tmX, tmY, q);
    }
    DDIMB_LAUNCH_CHECK();
    return DDIMB_OK;
  }
  int bw, bt;
  tile_128(W, &bw, &bt);
  TnParams p;
  tn_init(p);
  p.x_at = atom_of(Cin); p.y_at = atom_of(Cout);
  CUtensorMap tmX, tmY;
  int r = nhwc_map(&tmX, x, B, T, W, Cin, p.x_at, bw, bt);
  if (r) return r;
  r = nhwc_map(&tmY, dy, B, T, W, Cout, p.y_at, bw, bt);
  if (r) return r;
  nhwc_digits(p.x, bw, bt);
  nhwc_digits(p.y, bw, bt);
  p.nchunk[0] = ceil_div(W, bw); p.nchunk[1] = ceil_div(T, bt); p.nchunk[2] = B;
  p.stride_m = 9; p.stride_n = (int64_t)Cin * 9; p.out = dw;
  int nj = 0;
  for (int tap = 0; tap < 9; ++tap)
    for (int m0 = 0; m0 < Cin; m0 += 128) {
      TnJob& jb = p.jobs[nj++];
      jb.dx[0] = (int16_t)(tap % 3 - 1); jb.dx[1] = (int16_t)(tap / 3 - 1);
      jb.x_c0 = m0; jb.y_c0 = 0;
      jb.m_valid = Cin - m0 < 128 ? Cin - m0 : 128; jb.n_valid = Cout;
      jb.out_off = (int64_t)m0 * 9 + tap;
    }
  p.njobs = nj;
  return launch_tn(p, tmX, tmY, stream, 1);
}

// taps of a 4x4 stride-2 pad-1 kernel on the space-to-depth view: kernel column kx reads pixel pair (w' + pdx) half ph,
// kernel row ky reads row pair (t' + pdy) sub-row pr
static const int kPairShift[4] = {-1, 0, 0, 1}, kPairHalf[4] = {1, 0, 1, 0};

// csrc/wgrad_updown_wide.cuh: H = high-resolution 32-channel tensor [B][T][W][32], L = low-resolution 64-channel tensor [B][T/2][W/2][64],
// g = fp32 [64][32][4][4] (Downsample: dw[co][ci], Upsample: dw[ci][co]).
static int launch_wgrad_updown_wide(const void* h, const void* l, float* g, int B, int T, int W, cudaStream_t stream) {
  UdParams q;
  memset(&q, 0, sizeof(q));
  q.B = B; q.T = T; q.W = W; q.g = g;
  const int To = T / 2, Wo = W / 2;
  int rt = 64;
  while (rt > 4 && (long long)B * (Wo / 128) * ceil_div(To, rt) < 2LL * ddimb_num_sms()) rt /= 2;
  q.RT = rt; q.units_t = ceil_div(To, rt); q.num_units = B * (Wo / 128) * q.units_t;
  CUtensorMap tmH, tmL;
  {
    uint64_t dims[4] = {64, (uint64_t)Wo, (uint64_t)T, (uint64_t)B};                      // pixel pairs: 64 contiguous channels
    uint64_t str[3] = {64, (uint64_t)W * 32, (uint64_t)T * W * 32};
    uint32_t box[4] = {64, (uint32_t)kUdHPairs, 1, 1};
    int r = make_tmap_bf16(&tmH, h, 4, dims, str, box, 64);
    if (r) return r;
  }
  {
    uint64_t dims[4] = {64, (uint64_t)Wo, (uint64_t)To, (uint64_t)B};
    uint64_t str[3] = {64, (uint64_t)Wo * 64, (uint64_t)To * Wo * 64};
    uint32_t box[4] = {64, 128, 1, 1};
    int r = make_tmap_bf16(&tmL, l, 4, dims, str, box, 64);
    if (r) return r;
  }
  DDIMB_SMEM_ATTR(wgrad_updown_wide_kernel, kUdSmem);
  const int grid = q.num_units < ddimb_num_sms() ? q.num_units : ddimb_num_sms();
  wgrad_updown_wide_kernel<<<grid, kUdThreads, kUdSmem, stream>>>(tmH, tmL, q);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}

// ------------------------------------------------------------------------------------------------ Downsample weight gradient
// dw[co][ci][ky][kx] += sum dy[b,t',w',co] * x[b,2t'+ky-1,2w'+kx-1,ci]   (autograd of models/diffusion.py:70-78)
extern "C" int ddimb_conv_down4x4_wgrad(const void* x, const void* dy, float* dw, int B, int T, int W, int Cin, int Cout,
                                        cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && dy && dw && B >= 1 && T >= 2 && T % 2 == 0);
  const double npos = (double)B * (T / 2) * (W / 2);
  DDIMB_PROF(stream, "wgrad_down4x4", Cin, W, npos * (4.0 * Cin + Cout) * 2 + 64.0 * Cin * Cout, 32.0 * Cin * Cout * npos);
  DDIMB_CHECK_ARG(W >= 16 && W <= 256 && (W & (W - 1)) == 0 && Cin % 32 == 0 && Cout % 32 == 0 && Cin <= 256 && Cout <= 256);
  if (Cin == 32 && Cout == 64 && (W / 2) % 128 == 0) return launch_wgrad_updown_wide(x, dy, dw, B, T, W, stream);   // level 0 -> 1
  const int To = T / 2, Wo = W / 2;
  int bw, bt;
  tile_128(Wo, &bw, &bt);
  TnParams p;
  tn_init(p);
  p.x_at = atom_of(Cin); p.y_at = atom_of(Cout);
  CUtensorMap tmX, tmY;
  int r = s2d_map(&tmX, x, B, T, W, Cin, p.x_at, bw, bt);
  if (r) return r;
  r = nhwc_map(&tmY, dy, B, To, Wo, Cout, p.y_at, bw, bt);
  if (r) return r;
  s2d_digits(p.x, bw, bt);
  nhwc_digits(p.y, bw, bt);
  p.nchunk[0] = ceil_div(Wo, bw); p.nchunk[1] = ceil_div(To, bt); p.nchunk[2] = B;
  p.stride_m = 16; p.stride_n = (int64_t)Cin * 16; p.out = dw;
  int nj = 0;
  for (int ky = 0; ky < 4; ++ky)
    for (int kx = 0; kx < 4; ++kx)
      for (int m0 = 0; m0 < Cin; m0 += 128) {
        DDIMB_CHECK_ARG(nj < kTnMaxJobs);
        TnJob& jb = p.jobs[nj++];
        jb.dx[0] = (int16_t)kPairShift[kx]; jb.dx[1] = (int16_t)kPairHalf[ky]; jb.dx[2] = (int16_t)kPairShift[ky];
        jb.x_c0 = kPairHalf[kx] * Cin + m0; jb.y_c0 = 0;
        jb.m_valid = Cin - m0 < 128 ? Cin - m0 : 128; jb.n_valid = Cout;
        jb.out_off = (int64_t)m0 * 16 + ky * 4 + kx;
      }
  p.njobs = nj;
  return launch_tn(p, tmX, tmY, stream);
}

// ------------------------------------------------------------------------------------------------ Upsample weight gradient
// ConvTranspose2d weight [Cin][Cout][4][4]: dw[ci][co][ky][kx] += sum x[b,t,w,ci] * dy[b,2t+ky-1,2w+kx-1,co]
// (autograd of models/diffusion.py:59-67).  x [B][T][W][Cin], dy [B][2T][2W][Cout]; dy is the space-to-depth operand.
extern "C" int ddimb_conv_up4x4_wgrad(const void* x, const void* dy, float* dw, int B, int T, int W, int Cin, int Cout,
                                      cudaStream_t stream) {
  DDIMB_CHECK_ARG(x && dy && dw && B >= 1 && T >= 1);
  const double npos = (double)B * T * W;
  DDIMB_PROF(stream, "wgrad_up4x4", Cin, W, npos * (Cin + 4.0 * Cout) * 2 + 64.0 * Cin * Cout, 32.0 * Cin * Cout * npos);
  DDIMB_CHECK_ARG(W >= 8 && W <= 128 && (W & (W - 1)) == 0 && Cin % 32 == 0 && Cout % 32 == 0 && Cin <= 256 && Cout <= 256);
  if (Cin == 64 && Cout == 32 && W % 128 == 0) return launch_wgrad_updown_wide(dy, x, dw, B, 2 * T, 2 * W, stream);   // level 1 -> 0
  int bw, bt;
  tile_128(W, &bw, &bt);
  TnParams p;
  tn_init(p);
  p.x_at = atom_of(Cout); p.y_at = atom_of(Cin);
  CUtensorMap tmX, tmY;
  int r = s2d_map(&tmX, dy, B, 2 * T, 2 * W, Cout, p.x_at, bw, bt);
  if (r) return r;
  r = nhwc_map(&tmY, x, B, T, W, Cin, p.y_at, bw, bt);
  if (r) return r;
  s2d_digits(p.x, bw, bt);
  nhwc_digits(p.y, bw, bt);
  p.nchunk[0] = ceil_div(W, bw); p.nchunk[1] = ceil_div(T, bt); p.nchunk[2] = B;
  p.stride_m = 16; p.stride_n = (int64_t)Cout * 16; p.out = dw;
  int nj = 0;
  for (int ky = 0; ky < 4; ++ky)
    for (int kx = 0; kx < 4; ++kx)
      for (int m0 = 0; m0 < Cout; m0 += 128) {
        DDIMB_CHECK_ARG(nj < kTnMaxJobs);
        TnJob& jb = p.jobs[nj++];
        jb.dx[0] = (int16_t)kPairShift[kx]; jb.dx[1] = (int16_t)kPairHalf[ky]; jb.dx[2] = (int16_t)kPairShift[ky];
        jb.x_c0 = kPairHalf[kx] * Cout + m0; jb.y_c0 = 0;
        jb.m_valid = Cout - m0 < 128 ? Cout - m0 : 128; jb.n_valid = Cin;
        jb.out_off = (int64_t)m0 * 16 + ky * 4 + kx;
      }
  p.njobs = nj;
  return launch_tn(p, tmX, tmY, stream);
}

// ------------------------------------------------------------------------------------------------ nn.Linear weight gradient
// dw[n][k] += sum_r g[r][n] * a[r][k]; a bf16 [rows][K] (ld lda), g bf16 [rows][N] (ld ldg), dw fp32 [N][K].
extern "C" int ddimb_linear_wgrad(const void* a, long long lda, const void* g, long long ldg, float* dw, int rows, int N, int K,
                                  cudaStream_t stream) {
  DDIMB_CHECK_ARG(a && g && dw && rows >= 1);
  DDIMB_PROF(stream, "wgrad_linear", N, K, 2.0 * rows * (N + K) + 8.0 * N * K, 2.0 * rows * N * K);
  DDIMB_CHECK_ARG(K % 32 == 0 && N % 32 == 0 && lda % 8 == 0 && ldg % 8 == 0);
  TnParams p;
  tn_init(p);
  p.x_at = atom_of(K); p.y_at = atom_of(N);
  CUtensorMap tmX, tmY;
  {
    uint64_t dims[5] = {(uint64_t)K, (uint64_t)rows, 1, 1, 1};
    uint64_t str[4] = {(uint64_t)lda, (uint64_t)lda * rows, (uint64_t)lda * rows, (uint64_t)lda * rows};
    uint32_t box[5] = {(uint32_t)p.x_at, kTnChunk, 1, 1, 1};
    int r = make_tmap_bf16(&tmX, a, 5, dims, str, box, p.x_at);
    if (r) return r;
  }
  {
    uint64_t dims[5] = {(uint64_t)N, (uint64_t)rows, 1, 1, 1};
    uint64_t str[4] = {(uint64_t)ldg, (uint64_t)ldg * rows, (uint64_t)ldg * rows, (uint64_t)ldg * rows};
    uint32_t box[5] = {(uint32_t)p.y_at, kTnChunk, 1, 1, 1};
    int r = make_tmap_bf16(&tmY, g, 5, dims, str, box, p.y_at);
    if (r) return r;
  }
  p.x.dig[0] = 0; p.x.step[0] = kTnChunk;
  p.y.dig[0] = 0; p.y.step[0] = kTnChunk;
  p.nchunk[0] = ceil_div(rows, kTnChunk);
  p.stride_m = 1; p.stride_n = K; p.out = dw;
  // more (M tile, N tile) pairs than one launch's job table holds are issued as several launches
  int nj = 0;
  for (int n0 = 0; n0 < N; n0 += 256)
    for (int m0 = 0; m0 < K; m0 += 128) {
      TnJob& jb = p.jobs[nj++];
      memset(&jb, 0, sizeof(jb));
      jb.x_c0 = m0; jb.y_c0 = n0;
      jb.m_valid = K - m0 < 128 ? K - m0 : 128; jb.n_valid = N - n0 < 256 ? N - n0 : 256;
      jb.out_off = (int64_t)n0 * K + m0;
      if (nj == kTnMaxJobs) {
        p.njobs = nj;
        int r = launch_tn(p, tmX, tmY, stream);
        if (r) return r;
        nj = 0;
      }
    }
  if (nj) {
    p.njobs = nj;
    return launch_tn(p, tmX, tmY, stream);
  }
  return DDIMB_OK;
}
