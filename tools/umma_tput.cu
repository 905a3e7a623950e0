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
  int delay;      // dependent ALU instructions the issuing thread executes between iterations (stand-in for per-row bookkeeping)
  int issuers;    // warps issuing concurrently, each its own MMA stream into its own 128 accumulator columns (0 means 1)
  int shared_acc; // 1: all issuers accumulate into the SAME columns; operands are all ones, the sum is checked (lost updates = hazard)
};

__global__ void __launch_bounds__(128, 1) tput(Cfg c, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint32_t tmem_ptr;
  __shared__ __align__(8) uint64_t bar4[4];
  __shared__ __align__(8) uint64_t bar2[2];
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar4[i]), 1); mbar_init(smem_u32(&bar2[0]), 1); mbar_init(smem_u32(&bar2[1]), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_ptr), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  const int nis = c.issuers ? c.issuers : 1;
  if (c.shared_acc) {
    for (uint32_t i = threadIdx.x; i < 196u * 1024u / 4u; i += blockDim.x) reinterpret_cast<uint32_t*>(raw + (base - smem_u32(raw)))[i] = 0x3F803F80u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {                      // start the accumulator from one product (overwrite) and account for it in the check
      umma_bf16(tm, make_kmajor_desc(base, 8 * c.row_bytes, c.row_bytes == 128 ? 2 : 4), make_kmajor_desc(base + 96 * 1024, 8 * c.row_bytes, c.row_bytes == 128 ? 2 : 4), make_idesc_bf16(128, c.n), 0u);
      umma_commit(smem_u32(&bar2[0]));
      mbar_wait(smem_u32(&bar2[0]), 0, 98);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) < nis) {
    const uint32_t wq = threadIdx.x >> 5;
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
    uint32_t dly = 1;
    for (int it = 0; it < iters; ++it) {
      for (int q = 0; q < c.delay; ++q) asm volatile("mad.lo.u32 %0, %0, %0, %1;" : "+r"(dly) : "r"(q));
      const uint32_t t = tm + (c.shared_acc ? 0u : wq * 128) + (uint32_t)((it * c.col_step) & (nis > 1 ? 31 : 255));
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
    umma_commit(smem_u32(&bar4[wq]));
    mbar_wait(smem_u32(&bar4[wq]), 0, 99);
    long long t1 = clock64();
    if (blockIdx.x == 0) atomicMax((unsigned long long*)out, (unsigned long long)(t1 - t0 + (dly == 0x7fffffffu ? 1 : 0)));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (c.shared_acc && threadIdx.x < 32) {
    uint32_t v[16];
    tmem_ld16(tm, v);
    tmem_ld_wait();
    const float want = 16.f * (1.f + 6.f * iters * nis);
    int bad = 0;
    for (int i = 0; i < 16; ++i) bad += (__uint_as_float(v[i]) != want);
    if (bad) atomicAdd((unsigned long long*)(out + 1), 1ull);
    if (blockIdx.x == 0 && threadIdx.x == 0) out[2] = (long long)__uint_as_float(v[0]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d;
  cudaMalloc(&d, 24);
  cudaFuncSetAttribute(tput, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const Cfg cfgs[] = {
      {96, 64, 0, 2, 6, 0, 0, 0, 0},   {96, 64, 0, 2, 6, 0, 32, 1, 1}, {96, 64, 0, 2, 6, 0, 32, 2, 1},  {96, 64, 0, 2, 6, 0, 0, 2, 0},
      {192, 128, 0, 4, 12, 0, 0, 0, 0}, {192, 128, 0, 4, 12, 0, 64, 1, 1}, {192, 128, 0, 4, 12, 0, 64, 2, 1},
      // {n, row bytes, shift, kper, nmma, a_stride, col_step, commits, split, delay, issuers, shared_acc}
      {48, 64, 0, 2, 6, 0, 0, 0, 0, 0}, {48, 64, 0, 2, 6, 0, 0, 0, 0, 100}, {48, 64, 0, 2, 6, 0, 0, 2, 0, 0}, {96, 64, 0, 2, 6, 9216, 0, 0, 0, 100},
      {48, 64, 0, 2, 6, 9216, 0, 0, 0, 0, 2}, {48, 64, 0, 2, 6, 9216, 0, 0, 0, 0, 3}, {48, 64, 0, 2, 6, 9216, 0, 0, 0, 0, 4},
      {96, 64, 0, 2, 6, 9216, 0, 0, 0, 0, 2}, {96, 64, 0, 2, 6, 9216, 0, 0, 0, 0, 3}, {192, 128, 0, 4, 12, 0, 0, 0, 0, 0, 2},
      {48, 64, 0, 2, 6, 0, 0, 0, 0, 0, 2, 1}, {96, 64, 0, 2, 6, 0, 0, 0, 0, 0, 2, 1}, {96, 64, 0, 2, 6, 0, 0, 1, 0, 7, 3, 1}, {96, 64, 0, 2, 6, 0, 0, 0, 0, 3, 4, 1},
      {32, 64, 0, 2, 6, 0, 0, 0, 0},   {64, 64, 0, 2, 6, 0, 0, 0, 0},    {128, 64, 0, 2, 6, 0, 0, 0, 0}, {256, 64, 0, 2, 6, 0, 0, 0, 0},
  };
  const int iters = 2000;
  for (const Cfg& c : cfgs) {
    cudaMemset(d, 0, 24);
    tput<<<148, 128, 200 * 1024>>>(c, iters, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long res[3] = {0, 0, 0};
    cudaMemcpy(res, d, 24, cudaMemcpyDeviceToHost);
    const long long cyc = res[0];
    if (c.shared_acc) printf("  shared accumulator: %lld warps saw a wrong sum; value %lld, expected %lld\n", res[1], res[2], 16ll * (1 + 6ll * iters * (c.issuers ? c.issuers : 1)));
    const double per = (double)cyc / ((double)iters * 6);
    printf("N=%3d row=%3dB shift=%d a_stride=%5d col_step=%2d commits=%d split=%d delay=%3d issuers=%d : %.1f cycles/MMA-slot, %.0f / iteration (ideal %d)  %s\n", c.n, c.row_bytes, c.shift, c.a_stride, c.col_step, c.commits, c.split, c.delay, c.issuers ? c.issuers : 1, per, per * 6, c.n / 2,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
