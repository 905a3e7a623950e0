// Throughput probe: MUFU rates per SM for the activation candidates (tanh.approx.f32 / .f16x2 / .bf16x2, ex2, rcp).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/mufu_tput tools/mufu_tput.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP>
__device__ __forceinline__ uint32_t op(uint32_t x) {
  uint32_t y;
  if (OP == 0) asm volatile("tanh.approx.f32 %0, %1;" : "=r"(y) : "r"(x));
  if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=r"(y) : "r"(x));
  if (OP == 2) asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=r"(y) : "r"(x));
  if (OP == 3) asm volatile("tanh.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x));
  if (OP == 4) asm volatile("tanh.approx.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}

template <int OP>
__global__ void k(uint32_t* out, int iters, long long* cyc) {
  uint32_t v[8];
  for (int i = 0; i < 8; ++i) v[i] = threadIdx.x * 8 + i + 0x3c003c00u;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = op<OP>(v[i]);
  }
  __syncthreads();
  long long t1 = clock64();
  uint32_t s = 0;
  for (int i = 0; i < 8; ++i) s ^= v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP>
void run(const char* name, uint32_t* out, long long* cyc) {
  const int iters = 4096, threads = 512;
  k<OP><<<148, threads>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-20s %.2f results-ops/clk/SM (thread-instructions per cycle per SM)\n", name, (double)iters * 8 * threads / (double)c);
}

int main() {
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 512 * 4);
  cudaMalloc(&cyc, 8);
  run<0>("tanh.approx.f32", out, cyc);
  run<1>("ex2.approx.f32", out, cyc);
  run<2>("rcp.approx.f32", out, cyc);
  run<3>("tanh.approx.f16x2", out, cyc);
  run<4>("tanh.approx.bf16x2", out, cyc);
  return 0;
}
