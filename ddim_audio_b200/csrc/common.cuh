// Shared device/host helpers for libddim_b200 (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define DDIMB_OK 0
#define DDIMB_ERR_ARG (-1)      // bad argument / unsupported shape
#define DDIMB_ERR_DRIVER (-2)   // driver entry point (cuTensorMapEncodeTiled) unavailable
#define DDIMB_ERR_TMAP (-3)     // tensor-map encode failed

#define DDIMB_CHECK_ARG(cond)                                                                  \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      ddimb_set_error("%s:%d: argument check failed: %s", __FILE__, __LINE__, #cond);          \
      return DDIMB_ERR_ARG;                                                                    \
    }                                                                                          \
  } while (0)

#define DDIMB_CUDA(call)                                                                       \
  do {                                                                                         \
    cudaError_t e__ = (call);                                                                  \
    if (e__ != cudaSuccess) {                                                                  \
      ddimb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));  \
      return (int)e__;                                                                         \
    }                                                                                          \
  } while (0)

void ddimb_set_error(const char* fmt, ...);
// runtime.cu: launch counter and the optional per-launch CUDA-event profiler (bench.py's roofline leg)
void ddimb_count_launch();
struct DdimbProfScope {
  int slot;
  cudaStream_t stream;
  DdimbProfScope(cudaStream_t s, const char* name, int p0, int p1, double bytes, double flops);
  ~DdimbProfScope();
};
// name[p0,p1] identifies a kernel class; bytes / flops are the ALGORITHMIC figures of this one call
#define DDIMB_PROF(stream, name, p0, p1, bytes, flops) DdimbProfScope prof_scope__(stream, name, p0, p1, (double)(bytes), (double)(flops))

#define DDIMB_LAUNCH_CHECK()                                                                   \
  do {                                                                                         \
    ddimb_count_launch();                                                                      \
    cudaError_t e__ = cudaGetLastError();                                                      \
    if (e__ != cudaSuccess) {                                                                  \
      ddimb_set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return (int)e__;                                                                         \
    }                                                                                          \
  } while (0)

// per-device caches (a process may drive several GPUs): SM count, and "this kernel's dynamic shared-memory opt-in has been set"
static inline int ddimb_num_sms() {
  static int n[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (n[dev] == 0 && (cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n[dev] <= 0)) n[dev] = 148;
  return n[dev];
}
#define DDIMB_SMEM_ATTR(kernel, bytes)                                                                          \
  do {                                                                                                          \
    static bool done__[64] = {false};                                                                           \
    int dev__ = 0;                                                                                              \
    cudaGetDevice(&dev__);                                                                                      \
    if (dev__ < 0 || dev__ >= 64 || !done__[dev__]) {                                                           \
      DDIMB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));             \
      if (dev__ >= 0 && dev__ < 64) done__[dev__] = true;                                                       \
    }                                                                                                           \
  } while (0)

#ifdef __CUDACC__
namespace ddimb {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------- math
// silu(x) = x * sigmoid(x) = x * (0.5 * tanh(0.5 x) + 0.5): one MUFU (tanh.approx) instead of ex2 + rcp.
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float silu_fast(float x) { return x * fmaf(0.5f, tanh_approx(0.5f * x), 0.5f); }
// exact-ish variant for fp32 paths (temb MLP): expf + division
__device__ __forceinline__ float silu_exact(float x) { return x / (1.0f + __expf(-x)); }
// hf activations.py:59-66 NewGELUActivation
__device__ __forceinline__ float gelu_new_fast(float x) {
  const float k = 0.7978845608028654f;
  return 0.5f * x * (1.0f + tanh_approx(k * fmaf(0.044715f * x, x * x, x)));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// ---------------------------------------------------------------- packed fp32x2 math (sm_100: FFMA2 / FADD2, one issue slot for two lanes of work)
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
  return d;
}
__device__ __forceinline__ uint64_t pack2u(uint32_t lo, uint32_t hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
  return d;
}
__device__ __forceinline__ float2 unpack2(uint64_t v) {
  float2 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// two bf16 (one 32-bit word) -> fp32x2
__device__ __forceinline__ uint64_t bf16x2_to_f32x2(uint32_t w) { return pack2u(w << 16, w & 0xffff0000u); }
// fp32x2 -> two bf16 (round to nearest even)
__device__ __forceinline__ uint32_t f32x2_to_bf16x2(uint64_t v) {
  const float2 f = unpack2(v);
  return pack_bf16x2(f.x, f.y);
}
// SiLU of 2u given u = x/2:  x * sigmoid(x) = u * tanh(u) + u   (one MUFU per value, one FFMA2 per pair)
__device__ __forceinline__ uint64_t silu2_from_half(uint64_t u) {
  const float2 f = unpack2(u);
  return ffma2(u, pack2(tanh_approx(f.x), tanh_approx(f.y)), u);
}

// 128-bit streaming load / store
__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_v4(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// non-blocking probe (mbarrier.try_wait may suspend the thread for a while; this never does)
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap (loud failure) instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int tag) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3ff) == 0 && clock64() - t0 > 6000000000LL) {
      printf("ddimb: mbarrier timeout tag=%d block=%d thread=%d parity=%u\n", tag, blockIdx.x, threadIdx.x, parity);
      __trap();
    }
  }
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

// 1-D bulk copy global -> shared (no tensor map, no per-row cost), completion on an mbarrier
__device__ __forceinline__ void bulk_load_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
               "r"(bar)
               : "memory");
}

// ---- TMA stores (shared -> global).  Bulk async-groups are per THREAD: every lane of the warp executes the store wrapper (one
// elected lane issues), the commit and the wait, so whichever lane was elected also waits; for the others both are no-ops.
__device__ __forceinline__ void tma_store_4d_e(const void* tmap, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];\n\t}" ::"l"(tmap),
      "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// 1-D bulk store shared -> global (no tensor map, no per-box-row cost in the TMA engine); bytes a multiple of 16
__device__ __forceinline__ void bulk_store_s2g_e(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n\t}" ::"l"(dst),
      "r"(src), "r"(bytes)
      : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the shared-memory SOURCE of all but the newest N committed groups has been read (the buffer may be overwritten)
template <int N>
__device__ __forceinline__ void bulk_wait_group_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
// all but the newest N committed groups are complete (writes performed)
template <int N>
__device__ __forceinline__ void bulk_wait_group() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// ---------------------------------------------------------------- tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish() { asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem], bf16 inputs, fp32 accumulate, issued by ONE thread.
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
#ifdef DDIMB_NO_UMMA   // timing experiment: keep the issuing thread's bookkeeping, drop the tensor-core work
  asm volatile("" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate));
  return;
#endif
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once every tcgen05 op issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// One lane of a converged warp (CUTLASS-style elect_one_sync).  A region entered through this predicate is known to ptxas to
// be single-threaded, so it keeps tcgen05 operands in uniform registers; `if (lane == 0)` makes it wrap every tcgen05.mma in
// an ELECT / BRA.U.ANY waterfall loop.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "@p mov.s32 %0, 1;\n\t}"
      : "+r"(pred));
  return pred != 0;
}
// Warp-convergent variants: EVERY lane of the issuing warp executes these with identical (warp-uniform) operands and one
// elected lane issues.  With uniform control flow around them ptxas keeps the descriptors in uniform registers and emits a
// bare predicated UTCHMMA; the `if (lane == 0) { ... }` form makes it wrap every MMA in an ELECT / BRA.U.ANY waterfall loop
// (~150 issue cycles per MMA, measured: the issuing thread, not the tensor pipe, bounded the wide conv kernel).
__device__ __forceinline__ void umma_bf16_e(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_e(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar)
      : "memory");
}
// 32 lanes x 16 consecutive fp32 columns: thread i of the warp gets lane (base+i), columns [col, col+16)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// zero 32 lanes x 32 consecutive fp32 columns (thread i of the warp writes lane base + i)
__device__ __forceinline__ void tmem_st_zero32(uint32_t taddr) {
  asm volatile(
      "{\n\t.reg .b32 z;\n\tmov.b32 z, 0;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z};\n\t}" ::"r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_zero16(uint32_t taddr) {
  asm volatile(
      "{\n\t.reg .b32 z;\n\tmov.b32 z, 0;\n\t"
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {z,z,z,z,z,z,z,z,z,z,z,z,z,z,z,z};\n\t}" ::"r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// K-major shared-memory matrix descriptor (PTX ISA "tcgen05 shared memory descriptor"):
//   [0,14) start>>4 | [16,30) LBO>>4 (unused for swizzled K-major) | [32,46) SBO>>4 (8-row group pitch)
//   [46,48) version=1 | [49,52) base offset=0 (stage buffers are 1024-B aligned) | [61,64) swizzle (2=128B, 4=64B)
__device__ __forceinline__ uint64_t make_kmajor_desc(uint32_t smem_addr, uint32_t sbo_bytes, uint32_t swizzle_code) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)swizzle_code << 61;
  return d;
}
// Instruction descriptor, kind::f16: D=f32 (bits4-5=1), A=B=bf16 (bits7-9,10-12=1), K-major A/B, N>>3 @17, M>>4 @24
__device__ __forceinline__ uint32_t make_idesc_bf16(uint32_t m, uint32_t n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

}  // namespace ddimb
#endif  // __CUDACC__
