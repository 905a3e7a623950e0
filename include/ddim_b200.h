/* libddim_b200 -- C ABI of the B200-native DDIM-audio hot path.
 *
 * Drop-in boundary for klae01/ddim-audio: the reference has no FFI of its own (it is pure Python over torch.nn), so
 * these are the entry points its Python layer binds through ctypes (ddim_audio_b200/_C.py; INTEGRATION.md shows the
 * stub a maintainer would add to the reference).  Every function
 *   - takes raw DEVICE pointers, explicit sizes and a cudaStream_t as its last argument; no torch types;
 *   - is asynchronous on that stream, re-entrant, and owns no device memory (weight re-layouts live in caller-owned images,
 *     ddimb_conv_wimg);
 *   - returns 0 on success, a negative DDIMB_ERR_* code for argument/driver errors or a positive cudaError_t;
 *     ddimb_last_error() returns a thread-local description.
 * There is no CPU path: on a machine without an sm_100 device the launches fail with a CUDA error.
 *
 * Activation layout inside the network: NHWC bf16, [B][T][W][C] (the reference is NCHW fp32; conv_in / conv_out
 * convert at the network boundary).  GroupNorm statistics: double [B][8][2] = (sum, sum of squares) per group,
 * ACCUMULATED by the producing kernel into a buffer the caller has zeroed.
 */
#ifndef DDIM_B200_H_
#define DDIM_B200_H_

#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

int ddimb_version(void);
const char* ddimb_last_error(void);
/* Number of kernels this library has launched since it was loaded (bench.py's gpu_launches). */
long long ddimb_launch_count(void);
/* Opt-in profiler: while enabled every entry point below brackets its launches with two CUDA events on the caller's
 * stream.  ddimb_prof_report synchronises the device, writes a JSON array of per-kernel-class totals
 * {name, calls, launches, ms, bytes, flops} (bytes / flops = the algorithmic figures of DESIGN.md) into buf, clears the
 * records and returns the text length (negative on error). */
int ddimb_prof_enable(int on);
int ddimb_prof_report(char* buf, int buflen);

/* ---- sampler / loss arithmetic (fp32, elementwise) ------------------------------------------------------------ */

/* functions/denoising.py:27,41-43 -- one DDIM update, in place on xt:
 *   x0 = (xt - sqrt(1-at) et) / sqrt(at);  xt = sqrt(at_next) x0 + c2 et + c1 z.
 * z may be NULL (or c1 == 0): the noise term is skipped.  x0_out / xs_out (nullable) receive the snapshots of
 * :29-34 / :45-50.  n = number of floats (multiple of 4). */
int ddimb_ddim_step(float* xt, const float* et, const float* z, float* x0_out, float* xs_out, double at, double at_next,
                    double c1, double c2, long long n, cudaStream_t stream);

/* functions/losses.py:12-13 -- out = x0 sqrt(a[t]) + e sqrt(1 - a[t]); t is int64 [B]. */
int ddimb_qsample(const float* x0, const float* e, const float* alphas, const long long* t, float* out, int B,
                  long long per_sample, cudaStream_t stream);

/* functions/losses.py:15-18 -- per_sample[b] = sum (e - output)^2 over one sample (double; zeroed here). */
int ddimb_sqerr_loss(const float* e, const float* output, double* per_sample, int B, long long per_sample_elems,
                     cudaStream_t stream);

/* backward of ddimb_sqerr_loss: grad_output = -2 (e - output) g[b]; g is fp32 [B]. */
int ddimb_sqerr_loss_bwd(const float* e, const float* output, const float* g, float* grad_output, int B,
                         long long per_sample_elems, cudaStream_t stream);

/* ---- post-filter on the returned x0 snapshots --------------------------------------- */

/* External/UPU/signal/denoise.py:45-70 (denoise_2d, torch branch; called at runners/diffusion.py:400-401) on
 * x[B][C][T][F] fp32 (leading dims flattened into B): r = x - median_F(x); r -= median_T(r) (torch's LOWER median, exact);
 * then `iters` (the reference uses 5) times: denom = 1/(1e-9 + ||r||_2 over C), num = r denom,
 * r -= mean_F(num)/mean_F(denom), r -= mean_T(num)/mean_T(denom).  out may alias x.  C must be 2 (audio.yml) and
 * F one of 128 / 256 / 512, iters <= 8.  workspace: DEVICE scratch of at least ddimb_denoise2d_workspace(B, C, T, F, iters)
 * bytes (256-byte aligned; -1 = unsupported shape). */
long long ddimb_denoise2d_workspace(int B, int C, int T, int F, int iters);
int ddimb_denoise2d(const float* x, float* out, void* workspace, long long workspace_bytes, int B, int C, int T, int F, int iters,
                    cudaStream_t stream);

/* Allocation-free variant of ddimb_denoise2d.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same arguments, kernels and results as ddimb_denoise2d, but the scratch is a file-scope device array inside the
 * library (ddimb_denoise2d_static_max_bytes() = 4 MiB, held by every GPU that loads the library until exit) and the call performs no
 * cudaMalloc / cudaFree or other allocating call.  Shapes whose ddimb_denoise2d_workspace() exceeds that limit are rejected with
 * DDIMB_ERR_ARG and a message in ddimb_last_error() (x[3,2,8192,256] at 5 iterations fits, x[4,...] does not).  The array is shared
 * by every call on the GPU and keeps the last call's contents (what must start at zero is cleared on the stream at the start of each
 * call): calls that OVERLAP on the GPU -- different streams without ordering -- race on it; calls ordered on one stream do not. */
long long ddimb_denoise2d_static_max_bytes(void);
int ddimb_denoise2d_static(const float* x, float* out, int B, int C, int T, int F, int iters, cudaStream_t stream);

/* External/SST/utils/wav2img.py:82-127 pfft2wav as called by runners/diffusion.py:401-420 on x0.permute(0,3,2,1)[b], for a whole batch:
 * x fp32 [B][2][T][F] (NCHW: re / im planes) -> out int32 [B][(F-1)*(T-1)] PCM.  sri: double [F] ISO-226 weights of the rfft
 * frequencies; wt: fp32 [2F][NP] irfft basis (n = 2(F-1), NP = n rounded up to 128) with the hann synthesis window and win.sum()
 * folded in; norm: fp32 [n/2] overlap-add normaliser (all three are host-built constants, ddim_audio_b200/signal/pfft2wav.py).
 * hpi: the reference's HPI flag; target_norm, base_db, max_db, p_low, p_upp: wav2img.py:86-100.  ws: ddimb_pfft2wav_workspace() bytes. */
long long ddimb_pfft2wav_workspace(int B, int T, int F);
int ddimb_pfft2wav(const float* x, const double* sri, const float* wt, const float* norm, int* out, void* ws, long long ws_bytes, int B, int T,
                   int F, int hpi, double target_norm, double base_db, double max_db, double p_low, double p_upp, cudaStream_t stream);

/* ---- bf16 operand copies of the fp32 master weights, rebuilt in one launch (training: the masters change every step) ---------------- */

/* models/diffusion.py:28-78 weights in the operand layouts of the entry points below (tap-major 3x3 / 4x4 kernels, their flipped and
 * transposed data-gradient twins, Linear weights and transposes): each is a strided gather dst[i0][i1][i2][i3] = src[i0 s0 + i1 s1 +
 * i2 s2 + i3 s3] (negative strides = flips) with a cast to bf16.  chunk_table: DEVICE array of nchunks 80-byte records
 * {const float* src; void* dst; long long start, n; int d1, d2, d3, dst_f32; long long s0, s1, s2, s3} -- dst elements [start, start+n)
 * of a contiguous [d0][d1][d2][d3] copy (dst_f32: keep fp32); one CTA per record.  total_elems: profiler bookkeeping. */
int ddimb_repack(const void* chunk_table, int nchunks, long long total_elems, cudaStream_t stream);

/* ---- EMA of the parameters ------------------------------------------------------------- */

/* models/ema.py:16-23 -- shadow = (1 - mu) param + mu shadow for EVERY registered parameter in one launch (bit-exact: the
 * reference's operation order and fp32 roundings).  chunk_table: DEVICE array of nchunks records
 * {const float* param; float* shadow; long long count; long long pad} (32 bytes each, count <= 2^31; 128-bit path when both
 * pointers are 16-byte aligned),
 * one CTA per record; total_elems = sum of the counts (profiler bookkeeping only). */
int ddimb_ema_update(const void* chunk_table, int nchunks, long long total_elems, double mu, cudaStream_t stream);

/* ---- optimizer tail of the train step ------------------------------------------------ */

/* chunk_table for both calls: DEVICE array of nchunks 64-byte records
 *   {float* param; float* grad; float* exp_avg; float* exp_avg_var_or_sq; long long count; long long pad[3]}
 * (count <= 2^31; records whose pointers are all 16-byte aligned take the 128-bit path), one CTA per record; total_elems = sum
 * of counts (profiler bookkeeping).
 *
 * ddimb_grad_clip -- torch.nn.utils.clip_grad_norm_ as called at runners/diffusion.py:155-160: total_norm = ||all grads||_2
 * (fp64 accumulation), grads *= min(1, max_norm / (total_norm + 1e-6)) in place; total_norm_out (nullable, device float)
 * receives the norm.  sqnorm_scratch: one device double.
 *
 * ddimb_optim_step -- one optimizer step of every tensor in the table.  kind 0: torch.optim.AdamW (functions/__init__.py:15-23);
 * kind 1: AdaBelief as configured at functions/__init__.py:24-42 (AdaBelief.py:112-142: weight_decouple, no fixed_decay, no
 * amsgrad, no rectify, clip_step None).  step = the 1-based step count used for the bias corrections. */
int ddimb_grad_clip(const void* chunk_table, int nchunks, long long total_elems, double max_norm, double* sqnorm_scratch,
                    float* total_norm_out, cudaStream_t stream);
int ddimb_optim_step(const void* chunk_table, int nchunks, long long total_elems, int kind, double lr, double beta1, double beta2,
                     double eps, double weight_decay, long long step, cudaStream_t stream);

/* ---- kernel-ready weight images -------------------------------------------------------------------------------- */

/* Every convolution below reads its weights from a CALLER-OWNED device buffer ("image") that holds the weight tiles in the
 * exact shared-memory (TMA-swizzle) layout of the kernel that runs for (kind, Cin, Cout, W) -- tiles reach shared memory by
 * 1-D bulk copies, and nothing is re-laid out per call.  Build the image once per weight version (the reference's
 * nn.Conv2d.weight changes only at optimizer.step / load_state_dict) and pass it as `wimg` next to `w`.
 * kind: 0 Conv2d 3x3 (w bf16 [Cout][9*Cin]); 1 Conv2d 4x4 s2 (w bf16 [Cout][16*Cin]); 2 ConvTranspose2d 4x4 s2 (w bf16
 * [16*Cout][Cin]); 3 conv_in (w fp32 [32][2][3][3]); 4 conv_out (w fp32 [2][32][3][3]).  W: width of the op's INPUT tensor.
 * ddimb_conv_wimg_bytes: image size in bytes; 0 = this shape reads `w` directly (pass wimg NULL); < 0 = unsupported shape.
 * wimg must be 16-byte aligned. */
long long ddimb_conv_wimg_bytes(int kind, int Cin, int Cout, int W);
int ddimb_conv_wimg(int kind, const void* w, void* wimg, int Cin, int Cout, int W, cudaStream_t stream);
/* The same for n convolutions in one call: records = HOST array of n {int kind, Cin, Cout, W; const void* w; void* img} (32 bytes each). */
int ddimb_conv_wimg_batch(const void* records, int n, cudaStream_t stream);

/* ---- network boundary convolutions ------------------------------------------------------------------------------ */

/* models/diffusion.py:189-198 -- Conv2d(Cin=2 -> 32, k3 p1): x fp32 NCHW, w fp32 [32][Cin][3][3] (+ its image, kind 3, W = F),
 * out bf16 NHWC. */
int ddimb_conv_in(const float* x, const float* w, const void* wimg, const float* bias, void* out, double* stats, int B, int Cin, int T,
                  int F, int Cout, cudaStream_t stream);
/* models/diffusion.py:284,199-208 -- Conv2d(32 -> 2, k3 p1) of (x + skip): x, skip (nullable) bf16 NHWC,
 * w fp32 [2][32][3][3] (+ its image, kind 4, W = F), out fp32 NCHW. */
int ddimb_conv_out(const void* x, const void* skip, const float* w, const void* wimg, const float* bias, float* out, int B, int T, int F,
                   int Cin, int Cout, cudaStream_t stream);

/* ---- tcgen05 / TMEM / TMA implicit-GEMM convolutions and GEMMs ------------------------------------------------ */

/* models/diffusion.py:28-40,49,52 -- Conv2d 3x3 s1 p1 on NHWC bf16.  w: bf16 [Cout][9*Cin], k = (ky*3+kx)*Cin + ci.
 * Epilogue: + bias[Cout] (nullable) + bias_b[b*bias_b_stride + co] (nullable; the temb slice of :49), act (0 none,
 * 1 SiLU of :50,53, 3 training: store the pre-activation but take the statistics of its SiLU), + add (nullable bf16, same shape as out), GroupNorm partial sums into stats (nullable). */
int ddimb_conv3x3(const void* x, const void* w, const void* wimg, const float* bias, const float* bias_b, int bias_b_stride, int act,
                  const void* add, void* out, double* stats, int B, int T, int W, int Cin, int Cout, cudaStream_t stream);

/* models/diffusion.py:42-56 -- GroupNorm(8, C, eps) [+ SiLU] of x from precomputed statistics gn_stats [B][8][2], then the
 * Conv2d 3x3 of ddimb_conv3x3 with the same epilogue, in ONE pass over x (no materialised activation).  gn_mode 1: affine
 * only (norm[1] -> conv[1]); 2: affine + SiLU (norm[0], silu -> conv[0]).  gn_beta nullable.  Available for the shapes
 * ddimb_conv3x3_gn_supported() reports: C in {32, 64} with W a multiple of 128 (row-walking kernel, reads w directly) and C in
 * {96, 128} with 32 <= W <= 254 (window kernel, needs wimg; add must be NULL); other levels compose ddimb_gn_apply + ddimb_conv3x3. */
int ddimb_conv3x3_wide_supported(int Cin, int Cout, int W);
int ddimb_conv3x3_gn_supported(int C, int W);
int ddimb_conv3x3_gn(const void* x, const double* gn_stats, const float* gn_gamma, const float* gn_beta, float gn_eps, int gn_mode,
                     const void* w, const void* wimg, const float* bias, const float* bias_b, int bias_b_stride, int act, const void* add,
                     void* out, double* stats, int B, int T, int W, int C, cudaStream_t stream);

/* models/diffusion.py:70-78 -- Downsample Conv2d 4x4 s2 p1.  w: bf16 [Cout][16*Cin], k = (ky*4+kx)*Cin + ci.
 * x [B][T][W][Cin] -> out [B][T/2][W/2][Cout]. */
int ddimb_conv_down4x4(const void* x, const void* w, const void* wimg, const float* bias, void* out, double* stats, int B, int T, int W,
                       int Cin, int Cout, cudaStream_t stream);

/* models/diffusion.py:59-67 + :284 -- Upsample ConvTranspose2d 4x4 s2 p1, then "+ skip".
 * w: bf16 [16*Cout][Cin], row = (ky*4+kx)*Cout + co.  x [B][T][W][Cin] -> out [B][2T][2W][Cout]; skip nullable. */
int ddimb_conv_up4x4(const void* x, const void* w, const void* wimg, const float* bias, const void* skip, void* out, double* stats, int B,
                     int T, int W, int Cin, int Cout, cudaStream_t stream);

/* nn.Linear (models/diffusion.py:127-128,141,156,166; transformers modeling_fnet.py:204-231):
 * out[M][N] = act(a[M][K] w[N][K]^T + bias); a, w bf16 with leading dimensions lda, ldw; act 0 none / 1 SiLU /
 * 2 gelu_new; out bf16 or fp32 (out_f32) with leading dimension ldo; out_lo_offset != 0 also stores the bf16
 * rounding residual at out + out_lo_offset (split-bf16 operand for ddimb_fourier_mix). */
int ddimb_linear(const void* a, long long lda, const void* w, long long ldw, const float* bias, int act, void* out,
                 long long ldo, int out_f32, long long out_lo_offset, int M, int N, int K, cudaStream_t stream);

/* transformers modeling_fnet.py:142-177 -- y = Re(fftn(x, dim=(1,2))) for x [B][S][H] as two DFT-matrix GEMMs.
 * Operand layouts are documented at the definition (csrc/gemm_host.cu); split = 1 (bf16) or 3 (split-bf16). */
int ddimb_fourier_mix(const void* x, long long ldx, const void* wh, const void* a2, void* ut, float* y, int B, int S, int H,
                      int split, cudaStream_t stream);

/* ---- GroupNorm / LayerNorm / token repacks -------------------------------------------------------------------- */

/* models/diffusion.py:17-26,47-56 -- GroupNorm(8, C, eps) applied from precomputed statistics.
 * mode 0: y = gn(x); mode 1: y = silu(gn(x)); mode 2: y = res (+ res2) + gn(x) and accumulate stats_out of y;
 * modes 3 / 4 (training, x is a stored pre-activation): as 0 / 2 on silu(x).
 * beta nullable (norm[2] has no bias, :26).  res2 (nullable, mode 2): a second residual -- the last block of the network also
 * adds the skip operand of conv_out (:284), which then reads one tensor instead of two. */
int ddimb_gn_apply(const void* x, const void* res, const void* res2, void* y, const double* stats_in, double* stats_out, const float* gamma,
                   const float* beta, int mode, int B, long long positions, int C, float eps, cudaStream_t stream);

/* models/diffusion.py:107-120 -- BetaEmbedding layers in fp32: out[r][n] = act(in[idx?idx[r]:r] . w[n] + bias[n]);
 * act bit 0: SiLU on the output; bit 1: SiLU on the input (training keeps the pre-activations).  Any number of rows (chunks of 64). */
int ddimb_linear_small(const float* in, const long long* idx, const float* w, const float* bias, float* out, int rows,
                       int K, int N, int act, cudaStream_t stream);
/* out[r][:] = table[idx[r]][:] (fp32 rows of `width` values, width % 4 == 0): lookup into the timestep-embedding table
 * [num_timesteps][sum(embedding_size)] that eval-mode models precompute once per weight version (models/diffusion.py:107-120,248). */
int ddimb_gather_rows(const float* table, const long long* idx, float* out, int rows, int width, int table_rows, cudaStream_t stream);

/* models/diffusion.py:273-275,130-142 -- NHWC [B][S][W][C] -> tokens [B*S][C*W] (+ positional table te) -> LayerNorm. */
int ddimb_tokens_ln(const void* x, const float* te, const float* gamma, const float* beta, void* out, int B, int S, int W,
                    int C, float eps, cudaStream_t stream);
/* models/diffusion.py:277-279,284 -- tokens fp32 [B*S][C*W] -> NHWC bf16, + skip, GroupNorm partial sums. */
int ddimb_tokens_to_nhwc(const float* tok, const void* skip, void* out, double* stats, int B, int S, int W, int C,
                         cudaStream_t stream);
/* transformers modeling_fnet.py:180-188,226-231 -- y = LayerNorm_512(a + b) (b nullable); writes fp32 y and its bf16
 * operand copy ybf [rows][ldy] (columns [0,512) hi; [512,1024) rounding residual when write_lo). */
int ddimb_layernorm512(const float* a, const float* b, const float* gamma, const float* beta, float* y, void* ybf,
                       long long ldy, int write_lo, int rows, float eps, cudaStream_t stream);
/* fp32 [rows][H] -> bf16 hi (+ lo) planes of a [rows][ldy] operand buffer. */
int ddimb_split_bf16(const float* a, void* ybf, long long ldy, int H, int write_lo, long long rows, cudaStream_t stream);


/* ---- training: weight gradients (tcgen05, both operands MN-major; csrc/gemm_tn.cuh) ------------------------------ */
/* Every *_wgrad ACCUMULATES into a caller-zeroed fp32 gradient laid out exactly like the torch parameter. */

/* autograd of Conv2d 3x3 (models/diffusion.py:28-40): dw[Cout][Cin][3][3] += sum dy[b,t,w,co] x[b,t+ky-1,w+kx-1,ci]. */
int ddimb_conv3x3_wgrad(const void* x, const void* dy, float* dw, int B, int T, int W, int Cin, int Cout, cudaStream_t stream);
/* autograd of Downsample (models/diffusion.py:70-78): x [B][T][W][Cin], dy [B][T/2][W/2][Cout], dw[Cout][Cin][4][4]. */
int ddimb_conv_down4x4_wgrad(const void* x, const void* dy, float* dw, int B, int T, int W, int Cin, int Cout, cudaStream_t stream);
/* autograd of Upsample (models/diffusion.py:59-67): x [B][T][W][Cin], dy [B][2T][2W][Cout], dw[Cin][Cout][4][4]. */
int ddimb_conv_up4x4_wgrad(const void* x, const void* dy, float* dw, int B, int T, int W, int Cin, int Cout, cudaStream_t stream);
/* autograd of nn.Linear: dw[N][K] += g[rows][N]^T a[rows][K]; a, g bf16 with leading dimensions lda, ldg. */
int ddimb_linear_wgrad(const void* a, long long lda, const void* g, long long ldg, float* dw, int rows, int N, int K, cudaStream_t stream);

/* ---- training: GroupNorm / SiLU backward (models/diffusion.py:42-56) ------------------------------------------------ */
/* type 0: y = silu(gn(x)) (norm[0]); type 1: y = gn(silu(u)) (norm[1], norm[2]; u is the stored pre-activation).
 * g = dL/dy (bf16), z = x or u (bf16), stats = forward (sum, sumsq) of the GroupNorm input.
 * reduce: R[B][C][2] (double, caller-zeroed) = per-(sample, channel) sums the apply pass and the parameter gradients need.
 * apply:  out = dL/dx (+ add) for type 0, dL/du for type 1; sums_out (nullable, double [B][C], caller-zeroed) += sum over
 *         positions of out (the conv bias / temb-slice gradient).
 * param:  dgamma[C] (+= ) and dbeta[C] (+=, nullable). */
int ddimb_gn_bwd_reduce(const void* g, const void* z, const double* stats, const float* gamma, const float* beta, double* R, int type, int B,
                        long long positions, int C, float eps, cudaStream_t stream);
int ddimb_gn_bwd_apply(const void* g, const void* z, const void* add, void* out, const double* stats, const double* R, const float* gamma,
                       const float* beta, double* sums_out, int type, int B, long long positions, int C, float eps, cudaStream_t stream);
int ddimb_gn_bwd_param(const double* R, const double* stats, float* dgamma, float* dbeta, int B, long long positions, int C, float eps,
                       cudaStream_t stream);
/* out[c] += sum_r x[r][c] (bias gradients); x bf16 or fp32 [rows][ld]. */
int ddimb_colsum(const void* x, int is_bf16, long long ld, long long rows, int C, float* out, cudaStream_t stream);
/* double [B][C] sums -> out_b[b*ldb + c] += s (temb slice gradient, nullable) and out_c[c] += sum_b s (bias gradient, nullable). */
int ddimb_sums_scatter(const double* sums, float* out_b, long long ldb, float* out_c, int B, int C, cudaStream_t stream);

/* ---- training: FNet branch / token path / temb MLP / network boundary ------------------------------------------- */
/* backward of ddimb_layernorm512: dy = dy1 + dy2 (nullable); ds = dL/d(a+b) as fp32 and (dsbf nullable) bf16 hi | lo planes;
 * dgamma / dbeta accumulated. */
int ddimb_layernorm512_bwd(const float* a, const float* b, const float* dy1, const float* dy2, const float* gamma, float* ds, void* dsbf,
                           long long ldy, int write_lo, float* dgamma, float* dbeta, int rows, float eps, cudaStream_t stream);
/* hf activations.py:59-66 on a stored bf16 pre-activation: dact == NULL: out = gelu_new(u); else out = dact * gelu_new'(u). */
int ddimb_gelu(const void* u, const void* dact, void* out, long long n, cudaStream_t stream);
/* fp32 (a [+ b]) -> optional fp32 sum and bf16 hi (+ lo) operand planes [rows][ldy]. */
int ddimb_add_split_bf16(const float* a, const float* b, float* sum_out, void* ybf, long long ldy, int H, int write_lo, long long rows,
                         cudaStream_t stream);
/* nn.Dropout (models/diffusion.py:129,144; hf modeling_fnet.py:226-231): counter-based mask from (seed, index); the backward
 * is the same call on the gradient.  seed_dev (nullable, DEVICE): added to seed at run time -- a launch captured in a CUDA graph takes each
 * replay's seed from device memory. */
int ddimb_dropout(const float* x, float* y, long long n, float p, unsigned long long seed, const unsigned long long* seed_dev,
                  cudaStream_t stream);
/* backward of ddimb_tokens_ln (+ add: gradient of the additive skip reading the same tensor); dgamma / dbeta accumulated. */
int ddimb_tokens_ln_bwd(const void* x, const float* te, const float* gamma, const float* dtok, const void* add, void* dx, float* dgamma,
                        float* dbeta, int B, int S, int W, int C, float eps, cudaStream_t stream);
/* backward of ddimb_tokens_to_nhwc w.r.t. the tokens: NHWC bf16 [B][S][W][C] -> bf16 [B*S][C*W]. */
int ddimb_nhwc_to_tokens(const void* g, void* tok, int B, int S, int W, int C, cudaStream_t stream);
/* backward of ddimb_linear_small (act bit 0 = SiLU on the output, bit 1 = SiLU on the input):
 * wgrad: dw[N][K] += g^T f(in), db[N] += colsum(g); dgrad: out[r][k] = (z ? silu'(z[r][k]) : 1) sum_n g[r][n] w[n][k]. */
int ddimb_linear_small_wgrad(const float* g, const float* in, const long long* idx, float* dw, float* db, int rows, int K, int N, int in_act,
                             cudaStream_t stream);
int ddimb_linear_small_dgrad(const float* g, const float* w, const float* z, float* out, int rows, int K, int N, cudaStream_t stream);
/* parameter gradients of conv_in (g = g1 + g2, bf16 NHWC; g2 nullable) and conv_out (hs = h + skip). */
int ddimb_conv_in_wgrad(const float* x, const void* g1, const void* g2, float* dw, float* db, int B, int Cin, int T, int F, int Cout,
                        cudaStream_t stream);
int ddimb_conv_out_wgrad(const void* h, const void* skip, const float* dout, float* dw, float* db, int B, int T, int F, int Cin, int Cout,
                         cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DDIM_B200_H_ */
