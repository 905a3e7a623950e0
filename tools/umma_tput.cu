// Throughput probe: cycles per tcgen05.mma (M=128, K=16, bf16) as a function of N, the swizzle mode of the K-major
// operands and a row-shifted A start (what csrc/conv_wide.cuh issues).  Operand contents are irrelevant (garbage smem).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_tput tools/umma_tput.cu
#include <cstdio>
#include <cstdlib>
#include "../ddim_audio_b200/csrc/common.cuh"
using namespace ddimb;

struct Cfg {
  int n;          // UMMA N
  int row_bytes;  // 64 (SW64) or 128 (SW128)
  int shift;      // A start shifted by this many rows
  int kper;       // K=16 steps per (A,B) descriptor pair (row_bytes / 32)
  int nmma;       // MMAs per iteration
  int a_stride;   // bytes between the A tiles of consecutive MMAs groups (0: same tile every time)
  int col_step;   // accumulator column advance per iteration (the wide conv slides its N=3C window by C per input row)
  int commits;    // tcgen05.commit's (to a barrier nobody waits on) after each iteration's MMAs
  int split;      // 1: first MMA of an iteration is issued as N=2n/3 (accumulate) + N=n/3 (overwrite), like the wide conv's first step
};

__global__ void __launch_bounds__(128, 1) tput(Cfg c, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint32_t tmem_ptr;
  __shared__ __align__(8) uint64_t bar;
  __shared__ __align__(8) uint64_t bar2[2];
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_init(smem_u32(&bar2[0]), 1); mbar_init(smem_u32(&bar2[1]), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_ptr), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(128, c.n);
    const uint32_t sw = c.row_bytes == 128 ? 2 : 4;
    const uint32_t sbo = 8 * c.row_bytes;
    const uint32_t a0 = base + c.shift * c.row_bytes;
    const uint32_t b0 = base + 96 * 1024;
    // descriptors are built outside the timed loop: the loop body is MMA issue only (6 per iteration, fully unrolled)
    uint64_t ad[6], bd[6];
    for (int m = 0; m < 6; ++m) {
      const int grp = m / c.kper, k = m % c.kper;
      ad[m] = make_kmajor_desc(a0 + (grp % 4) * c.a_stride, sbo, sw) + 2 * k;
      bd[m] = make_kmajor_desc(b0 + (grp % 3) * c.n * c.row_bytes, sbo, sw) + 2 * k;
    }
    long long t0 = clock64();
    const uint32_t idesc_a = make_idesc_bf16(128, 2 * c.n / 3), idesc_b = make_idesc_bf16(128, c.n / 3);
    for (int it = 0; it < iters; ++it) {
      const uint32_t t = tm + (uint32_t)((it * c.col_step) & 255);
      if (c.split) {
        umma_bf16(t, ad[0], bd[0], idesc_a, 1u);
        umma_bf16(t + 2 * c.n / 3, ad[0], bd[0], idesc_b, 0u);
      } else {
        umma_bf16(t, ad[0], bd[0], idesc, 1u);
      }
#pragma unroll
      for (int m = 1; m < 6; ++m) umma_bf16(t, ad[m], bd[m], idesc, 1u);
      for (int q = 0; q < c.commits; ++q) umma_commit(smem_u32(&bar2[q]));
    }
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0, 99);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  cudaFuncSetAttribute(tput, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const Cfg cfgs[] = {
      {96, 64, 0, 2, 6, 0, 0, 0, 0},   {96, 64, 0, 2, 6, 0, 32, 1, 1}, {96, 64, 0, 2, 6, 0, 32, 2, 1},  {96, 64, 0, 2, 6, 0, 0, 2, 0},
      {192, 128, 0, 4, 12, 0, 0, 0, 0}, {192, 128, 0, 4, 12, 0, 64, 1, 1}, {192, 128, 0, 4, 12, 0, 64, 2, 1},
      {32, 64, 0, 2, 6, 0, 0, 0, 0},   {64, 64, 0, 2, 6, 0, 0, 0, 0},    {128, 64, 0, 2, 6, 0, 0, 0, 0}, {256, 64, 0, 2, 6, 0, 0, 0, 0},
  };
  const int iters = 2000;
  for (const Cfg& c : cfgs) {
    tput<<<148, 128, 200 * 1024>>>(c, iters, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long cyc = 0;
    cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
    const double per = (double)cyc / ((double)iters * 6);
    printf("N=%3d row=%3dB shift=%d a_stride=%5d col_step=%2d commits=%d split=%d : %.1f cycles/MMA-slot (ideal %d)  %s\n", c.n, c.row_bytes, c.shift, c.a_stride, c.col_step, c.commits, c.split, per, c.n / 2,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
