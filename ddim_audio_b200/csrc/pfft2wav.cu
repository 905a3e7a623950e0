// Output codec after the sampling loop: reference runners/diffusion.py:401-420 permutes every predicted
// x0 [B,2,T,F] to [B,F,T,2] and hands it sample by sample to External/SST/utils/wav2img.py:82-127 pfft2wav -- ISO-226 loudness
// statistic per frame, mean of its 92.5 .. 97.5 percentile window, gain = min(73, 86 - excess) dB, scipy.signal.istft with its defaults
// (irfft of n = 2 (F - 1) per frame x win.sum(), hann overlap-add with hop n / 2, n / 2 samples trimmed at both ends, division by the
// summed win^2), clip, int32.  Here the whole batch is processed on the device straight from the NCHW fp32 tensor (the permute is
// only an index map):
//   1. loud_rows_kernel     s[b][t] = sum_f |X[b][f][t]| (x sri[f] unless HPI)                      fp64 accumulate
//   2. window_gain_kernel   per sample: bitonic sort of s[b][:] in shared memory, mean of the rank window, gain[b]
//   3. istft_gemm_kernel    Y[b*T + t][m] = sum_k A[t][k] Wt[k][m]: the per-frame irfft as a real fp32 GEMM against the cos / sin
//                           basis with the synthesis window and win.sum() folded in (fp32 like the reference's own FFT; a [510 x 512]
//                           DFT per frame is 0.5 GFLOP per sample-second of work, nowhere near a bottleneck, so CUDA-core FMA keeps
//                           fp32-class accuracy without split-precision tensor-core tricks)
//   4. overlap_add_kernel   out[b][n] = int32(clip((Y[ii][m] + Y[ii-1][m + n/2]) / norm[m] * gain[b]))
// The float -> int conversion reproduces the reference's numpy semantics: np.clip against float32(2^31 - 1) = 2^31, then an x86
// truncating cast, which yields INT_MIN for +2^31 (positive saturation wraps; so does NaN).
#include <math.h>

#include "../../include/ddim_b200.h"
#include "common.cuh"

namespace {

constexpr int kMaxT = 16384;   // frames per sample the in-shared-memory sort holds (doubles: 128 KB)

__global__ void __launch_bounds__(256) loud_rows_kernel(const float* __restrict__ x, const double* __restrict__ sri, double* __restrict__ s, int T, int F,
                                                        int hpi) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);   // b * T + t
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.y;
  if (row >= T) return;
  const float* re = x + ((int64_t)b * 2 * T + row) * F;
  const float* im = re + (int64_t)T * F;
  double acc = 0.0;
  for (int f = lane; f < F; f += 32) {
    const float a = hypotf(re[f], im[f]);
    acc += hpi ? (double)a : (double)a * sri[f];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) s[(int64_t)b * T + row] = acc;
}

// one CTA per sample: ascending bitonic sort of T values padded to a power of two with +inf, then the window mean and the gain
__global__ void __launch_bounds__(1024) window_gain_kernel(const double* __restrict__ s, float* __restrict__ gain, int T, int npow2, int lo, int hi,
                                                           double target_norm, double base_db, double max_db) {
  extern __shared__ double sm[];
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < npow2; i += blockDim.x) sm[i] = i < T ? s[(int64_t)b * T + i] : INFINITY;
  __syncthreads();
  for (int k = 2; k <= npow2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const double a = sm[i], c = sm[l];
          const bool up = (i & k) == 0;
          if ((a > c) == up) { sm[i] = c; sm[l] = a; }
        }
      }
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    double acc = 0.0;
    for (int i = lo; i <= hi; ++i) acc += sm[i];
    const double cur = acc / (double)(hi - lo + 1);
    const double exp_amp = log10(cur / target_norm + 1e-9) * 10.0;     // wav2img.py:115
    const double amp = fmin(base_db, max_db - exp_amp);                 // :121
    gain[b] = (float)pow(10.0, amp / 10.0);                             // :123 (the Python float meets a float32 array: rounded to float32)
  }
}

// Y[M][NP] = A[M][K] Wt[K][NP]; A row r = (b, t): k < F -> re plane, k >= F -> im plane of x (each value divided by sri when HPI).
constexpr int BM = 128, BN = 128, BK = 16;
__global__ void __launch_bounds__(256) istft_gemm_kernel(const float* __restrict__ x, const double* __restrict__ sri, const float* __restrict__ wt,
                                                         float* __restrict__ y, int64_t M, int T, int F, int NP, int hpi) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.y * BM;
  const int n0 = blockIdx.x * BN;
  const int K = 2 * F;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += BK) {
    // A tile: 128 rows x 16 k as 512 float4 (k contiguous in memory inside a plane; BK divides F)
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256, r = idx >> 2, kq = (idx & 3) * 4;
      const int64_t row = m0 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < M) {
        const int64_t b = row / T, t = row - b * T;
        const int k = k0 + kq, plane = k >= F ? 1 : 0, f = k - plane * F;
        v = *reinterpret_cast<const float4*>(x + ((b * 2 + plane) * T + t) * F + f);
        if (hpi) {
          v.x = (float)((double)v.x / sri[f]); v.y = (float)((double)v.y / sri[f + 1]);
          v.z = (float)((double)v.z / sri[f + 2]); v.w = (float)((double)v.w / sri[f + 3]);
        }
      }
      As[kq][r] = v.x; As[kq + 1][r] = v.y; As[kq + 2][r] = v.z; As[kq + 3][r] = v.w;
    }
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256, kk = idx >> 5, nq = (idx & 31) * 4;
      *reinterpret_cast<float4*>(&Bs[kk][nq]) = *reinterpret_cast<const float4*>(wt + (int64_t)(k0 + kk) * NP + n0 + nq);
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[8], bb[8];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 8]), a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 8 + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 8]), b1 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 8 + 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      bb[0] = b0.x; bb[1] = b0.y; bb[2] = b0.z; bb[3] = b0.w; bb[4] = b1.x; bb[5] = b1.y; bb[6] = b1.z; bb[7] = b1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t row = m0 + ty * 8 + i;
    if (row < M) {
      float* yp = y + row * NP + n0 + tx * 8;
      *reinterpret_cast<float4*>(yp) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      *reinterpret_cast<float4*>(yp + 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
    }
  }
}

__global__ void __launch_bounds__(256) overlap_add_kernel(const float* __restrict__ y, const float* __restrict__ norm, const float* __restrict__ gain,
                                                          int* __restrict__ out, int T, int NP, int half, int64_t L) {
  const int b = blockIdx.y;
  const float g = gain[b];
  const float* yb = y + (int64_t)b * T * NP;
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < L; n += (int64_t)gridDim.x * blockDim.x) {
    const int64_t nn = n + half;                 // position before the boundary trim
    const int64_t ii = nn / half;                // 1 .. T - 1: segment ii starts at ii * half
    const int m = (int)(nn - ii * half);
    float v = yb[(ii - 1) * NP + m + half] + yb[ii * NP + m];   // segment ii - 1 was added first (istft's loop order)
    v = v / norm[m];
    v = v * g;
    v = fminf(fmaxf(v, -2147483648.0f), 2147483648.0f);       // np.clip(r, -2^31, float32(2^31 - 1) == 2^31)
    out[(int64_t)b * L + n] = v >= 2147483648.0f ? (int)0x80000000 : (int)v;   // x86 cvttss2si: out of range -> INT_MIN
  }
}

}  // namespace

static int next_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// bytes of workspace for ddimb_pfft2wav: loudness rows (fp64) | gains | the windowed irfft segments Y [B*T][NP]
extern "C" long long ddimb_pfft2wav_workspace(int B, int T, int F) {
  if (B < 1 || T < 2 || T > kMaxT || F < 16 || F % 16 != 0 || F > 1024) return -1;
  const long long NP = ((2LL * (F - 1) + 127) / 128) * 128;
  return (((long long)B * T * 8 + 255) / 256) * 256 + ((B * 4LL + 255) / 256) * 256 + (long long)B * T * NP * 4;
}

extern "C" int ddimb_pfft2wav(const float* x, const double* sri, const float* wt, const float* norm, int* out, void* ws, long long ws_bytes, int B,
                              int T, int F, int hpi, double target_norm, double base_db, double max_db, double p_low, double p_upp,
                              cudaStream_t stream) {
  const long long need = ddimb_pfft2wav_workspace(B, T, F);
  DDIMB_CHECK_ARG(need > 0 && x && sri && wt && norm && out && ws && ws_bytes >= need && B <= 65535);
  DDIMB_CHECK_ARG(p_low >= 0 && p_low <= p_upp && p_upp <= 1 && target_norm != 0);
  const int n = 2 * (F - 1), half = n / 2;
  const int NP = ((n + 127) / 128) * 128;
  const long long L = (long long)half * (T - 1);
  double* s = reinterpret_cast<double*>(ws);
  DDIMB_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 255) == 0);
  float* gain = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + (((long long)B * T * 8 + 255) / 256) * 256);
  float* y = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(gain) + ((B * 4LL + 255) / 256) * 256);
  DDIMB_PROF(stream, "pfft2wav", T, F, (double)B * T * F * 8 + (double)B * L * 4, 2.0 * B * T * (2.0 * F) * n);
  loud_rows_kernel<<<dim3((T + 7) / 8, B), 256, 0, stream>>>(x, sri, s, T, F, hpi);
  DDIMB_LAUNCH_CHECK();
  const int npow2 = next_pow2(T);
  const int lo = (int)(T * p_low), hi = (int)(T * p_upp);      // wav2img.py:112 int(len(s_HPI) * p)
  DDIMB_CHECK_ARG(hi < T && lo <= hi);
  const int smem = npow2 * 8;
  DDIMB_SMEM_ATTR(window_gain_kernel, kMaxT * 8);
  window_gain_kernel<<<B, 1024, smem, stream>>>(s, gain, T, npow2, lo, hi, target_norm, base_db, max_db);
  DDIMB_LAUNCH_CHECK();
  const int64_t M = (int64_t)B * T;
  istft_gemm_kernel<<<dim3(NP / BN, (unsigned)((M + BM - 1) / BM)), 256, 0, stream>>>(x, sri, wt, y, M, T, F, NP, hpi);
  DDIMB_LAUNCH_CHECK();
  overlap_add_kernel<<<dim3((unsigned)((L + 1023) / 1024 < 4096 ? (L + 1023) / 1024 : 4096), B), 256, 0, stream>>>(y, norm, gain, out, T, NP, half, L);
  DDIMB_LAUNCH_CHECK();
  return DDIMB_OK;
}
