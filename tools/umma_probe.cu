// Probe: which shared-memory layouts let a tcgen05.mma A-operand descriptor start at an arbitrary ROW (position)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe tools/umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "../ddim_audio_b200/csrc/common.cuh"
using namespace ddimb;

constexpr int P = 176;   // positions held in smem
constexpr int N = 32;

// mode 0: no-swizzle planes [C/8][P][16B]; mode 1: SW128 rows of 128 B (C=64); mode 2: SW64 rows of 64 B (C=32)
__global__ void probe(const __nv_bfloat16* A, const __nv_bfloat16* Wt, float* D, int C, int mode, int shift, int bo_mode) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  uint8_t* smA = sm;                    // up to 176*128 = 22528 B
  uint8_t* smB = sm + 24576;            // planes [C/8][N][16B]
  __shared__ uint32_t tmem_ptr;
  __shared__ __align__(8) uint64_t bar;
  for (int i = threadIdx.x; i < P * C; i += blockDim.x) {
    const int p = i / C, c = i % C, j = c / 8;
    uint32_t off;
    if (mode == 0) off = (j * P + p) * 16 + (c % 8) * 2;
    else if (mode == 1) off = p * 128 + ((j ^ (p & 7)) << 4) + (c % 8) * 2;
    else off = p * 64 + ((j ^ ((p >> 1) & 3)) << 4) + (c % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(smA + off) = A[i];
  }
  for (int i = threadIdx.x; i < N * C; i += blockDim.x) {
    const int n = i / C, c = i % C, j = c / 8;
    *reinterpret_cast<__nv_bfloat16*>(smB + (j * N + n) * 16 + (c % 8) * 2) = Wt[i];
  }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_ptr), 32); tmem_relinquish(); }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(128, N);
    for (int k = 0; k < C / 16; ++k) {
      uint64_t ad = 0, bd = 0;
      uint32_t astart;
      if (mode == 0) {
        astart = base + shift * 16 + k * 2 * P * 16;
        ad = (uint64_t)((astart & 0x3FFFF) >> 4) | ((uint64_t)((P * 16) >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
      } else if (mode == 1) {
        astart = base + shift * 128 + k * 32;
        ad = (uint64_t)((astart & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
        if (bo_mode == 1) ad |= (uint64_t)((astart >> 7) & 7) << 49;
      } else {
        astart = base + shift * 64 + k * 32;
        ad = (uint64_t)((astart & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | (1ull << 46) | (4ull << 61);
        if (bo_mode == 1) ad |= (uint64_t)((astart >> 7) & 7) << 49;
        if (bo_mode == 2) ad |= (uint64_t)((astart >> 7) & 3) << 49;
      }
      const uint32_t bstart = base + 24576 + k * 2 * N * 16;
      bd = (uint64_t)((bstart & 0x3FFFF) >> 4) | ((uint64_t)((N * 16) >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
      umma_bf16(tm, ad, bd, idesc, k > 0);
    }
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0, 99);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tm + ((uint32_t)(warp * 32) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) D[(warp * 32 + lane) * N + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 32);
}

int main() {
  const int modes[3] = {0, 1, 2}, Cs[3] = {64, 64, 32};
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int mi = 0; mi < 3; ++mi) {
    const int C = Cs[mi], mode = modes[mi];
    std::vector<__nv_bfloat16> hA(P * C), hW(N * C);
    std::vector<float> fA(P * C), fW(N * C);
    srand(1 + mi);
    for (int i = 0; i < P * C; ++i) { float v = (rand() % 2001 - 1000) / 500.f; hA[i] = __float2bfloat16(v); fA[i] = __bfloat162float(hA[i]); }
    for (int i = 0; i < N * C; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hW[i] = __float2bfloat16(v); fW[i] = __bfloat162float(hW[i]); }
    __nv_bfloat16 *dA, *dW; float* dD;
    cudaMalloc(&dA, P * C * 2); cudaMalloc(&dW, N * C * 2); cudaMalloc(&dD, 128 * N * 4);
    cudaMemcpy(dA, hA.data(), P * C * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dW, hW.data(), N * C * 2, cudaMemcpyHostToDevice);
    const int shifts[] = {0, 1, 2, 3, 7, 8, 9, 17, 40};
    for (int bo = 0; bo < (mode == 0 ? 1 : (mode == 1 ? 2 : 3)); ++bo)
      for (int s : shifts) {
        cudaMemset(dD, 0, 128 * N * 4);
        probe<<<1, 128, 65536>>>(dA, dW, dD, C, mode, s, bo);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d shift %d bo %d: CUDA error %s\n", mode, s, bo, cudaGetErrorString(e)); return 1; }
        std::vector<float> hD(128 * N);
        cudaMemcpy(hD.data(), dD, 128 * N * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0;
        for (int m = 0; m < 128; ++m)
          for (int n = 0; n < N; ++n) {
            double ref = 0;
            for (int c = 0; c < C; ++c) ref += (double)fA[(m + s) * C + c] * fW[n * C + c];
            maxerr = fmax(maxerr, fabs(ref - hD[m * N + n]));
          }
        printf("mode %d (C=%d) base_offset_mode %d shift %2d: max err %.4g %s\n", mode, C, bo, s, maxerr, maxerr < 1e-2 ? "OK" : "WRONG");
      }
  }
  return 0;
}
