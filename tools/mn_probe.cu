// Probe: tcgen05.mma with MN-major operands (the weight-gradient GEMM contracts over POSITIONS, so both operands are
// channel-contiguous = MN-major when they are the NHWC tiles TMA delivers).  D[m][n] = sum_p X[p][m] * Y[p][n].
// X atoms: [P rows][AT channels] swizzled tiles (AT = 64 for SWIZZLE_128B, 32 for SWIZZLE_64B) at a uniform byte stride (LBO).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/mn_probe tools/mn_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../ddim_audio_b200/csrc/common.cuh"
using namespace ddimb;

constexpr int P = 80;   // positions (K) held per atom tile; the MMAs use rows [shift, shift + KP)
constexpr int KP = 64;

struct Cfg {
  int sw;       // 128 or 64
  int m_atoms;  // atoms along M present in smem (M of the instruction is mshape)
  int n_atoms;
  int mshape;   // 128 or 64
  int shift;    // start row of X
};

__global__ void probe(const __nv_bfloat16* X, const __nv_bfloat16* Y, float* D, Cfg c) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* sm = raw + (base - smem_u32(raw));
  const int AT = c.sw == 128 ? 64 : 32;
  const int rowb = AT * 2;
  const int tile_bytes = ((P * rowb + 1023) / 1024) * 1024;
  const int CX = c.m_atoms * AT, CY = c.n_atoms * AT;
  uint8_t* smX = sm;
  uint8_t* smY = sm + 4 * 16384;
  __shared__ uint32_t tmem_ptr;
  __shared__ __align__(8) uint64_t bar;
  for (int i = threadIdx.x; i < 8 * 16384; i += blockDim.x) sm[i] = 0;
  __syncthreads();
  auto put = [&](uint8_t* dst, int p, int ch, __nv_bfloat16 v) {
    const int atom = ch / AT, cc = ch % AT, j = cc / 8;
    const int sw = c.sw == 128 ? (p & 7) : ((p >> 1) & 3);
    *reinterpret_cast<__nv_bfloat16*>(dst + atom * tile_bytes + p * rowb + ((j ^ sw) << 4) + (cc % 8) * 2) = v;
  };
  for (int i = threadIdx.x; i < P * CX; i += blockDim.x) put(smX, i / CX, i % CX, X[i]);
  for (int i = threadIdx.x; i < P * CY; i += blockDim.x) put(smY, i / CY, i % CY, Y[i]);
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_ptr), 256); tmem_relinquish(); }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  const int N = CY;
  if (threadIdx.x == 0) {
    // kind::f16, D=f32, A=B=bf16, a_major (bit 15) = b_major (bit 16) = 1 (MN-major)
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(c.mshape >> 4) << 24);
    const uint32_t sbo = 8 * rowb;   // next group of 8 k rows
    const uint64_t code = c.sw == 128 ? 2 : 4;
    for (int k = 0; k < KP / 16; ++k) {
      const uint32_t xs = base + (c.shift + k * 16) * rowb;
      const uint32_t ys = base + 4 * 16384 + (k * 16) * rowb;
      const uint64_t ad = (uint64_t)((xs & 0x3FFFF) >> 4) | ((uint64_t)(tile_bytes >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | (code << 61);
      const uint64_t bd = (uint64_t)((ys & 0x3FFFF) >> 4) | ((uint64_t)(tile_bytes >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | (code << 61);
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
    for (int j = 0; j < 16; ++j) D[(warp * 32 + lane) * 256 + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 256);
}

int main() {
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16384 + 2048);
  const Cfg cfgs[] = {{128, 2, 1, 128, 0}, {128, 2, 2, 128, 0}, {128, 2, 1, 128, 3}, {128, 2, 2, 128, 9}, {128, 1, 1, 128, 0}, {128, 1, 2, 64, 0},
                      {64, 4, 1, 128, 0},  {64, 4, 3, 128, 0},  {64, 4, 2, 128, 5}, {64, 3, 3, 128, 0}, {64, 2, 1, 64, 0}};
  for (const Cfg& c : cfgs) {
    const int AT = c.sw == 128 ? 64 : 32, CX = c.m_atoms * AT, CY = c.n_atoms * AT;
    std::vector<__nv_bfloat16> hX(P * CX), hY(P * CY);
    std::vector<float> fX(P * CX), fY(P * CY);
    srand(7);
    for (int i = 0; i < P * CX; ++i) { float v = (rand() % 2001 - 1000) / 500.f; hX[i] = __float2bfloat16(v); fX[i] = __bfloat162float(hX[i]); }
    for (int i = 0; i < P * CY; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hY[i] = __float2bfloat16(v); fY[i] = __bfloat162float(hY[i]); }
    __nv_bfloat16 *dX, *dY; float* dD;
    cudaMalloc(&dX, P * CX * 2); cudaMalloc(&dY, P * CY * 2); cudaMalloc(&dD, 128 * 256 * 4);
    cudaMemcpy(dX, hX.data(), P * CX * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dY, hY.data(), P * CY * 2, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, 128 * 256 * 4);
    probe<<<1, 128, 8 * 16384 + 2048>>>(dX, dY, dD, c);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("sw%d m_atoms %d n_atoms %d M%d shift %d: CUDA error %s\n", c.sw, c.m_atoms, c.n_atoms, c.mshape, c.shift, cudaGetErrorString(e)); return 1; }
    std::vector<float> hD(128 * 256);
    cudaMemcpy(hD.data(), dD, 128 * 256 * 4, cudaMemcpyDeviceToHost);
    // reference rows m < min(mshape, CX); find for each logical row the lane that holds it
    const int rows = c.mshape < CX ? c.mshape : CX;
    double maxerr = 0;
    int misplaced = 0;
    std::vector<int> lane_of(rows, -1);
    for (int m = 0; m < rows; ++m) {
      std::vector<double> ref(CY);
      for (int n = 0; n < CY; ++n) {
        double r = 0;
        for (int p = 0; p < KP; ++p) r += (double)fX[(p + c.shift) * CX + m] * fY[p * CY + n];
        ref[n] = r;
      }
      for (int l = 0; l < 128 && lane_of[m] < 0; ++l) {
        double err = 0;
        for (int n = 0; n < CY; ++n) err = fmax(err, fabs(ref[n] - hD[l * 256 + n]));
        if (err < 2e-2) lane_of[m] = l;
      }
      if (lane_of[m] != m) ++misplaced;
      double err = 0;
      for (int n = 0; n < CY; ++n) err = fmax(err, fabs(ref[n] - hD[m * 256 + n]));
      maxerr = fmax(maxerr, err);
    }
    printf("sw%d m_atoms %d n_atoms %d M%d shift %d: identity-lane max err %.4g %s", c.sw, c.m_atoms, c.n_atoms, c.mshape, c.shift, maxerr, maxerr < 2e-2 ? "OK" : "WRONG");
    if (misplaced) {
      printf("  lanes:");
      for (int m = 0; m < rows; m += 8) printf(" %d->%d", m, lane_of[m]);
    }
    printf("\n");
    cudaFree(dX); cudaFree(dY); cudaFree(dD);
  }
  return 0;
}
